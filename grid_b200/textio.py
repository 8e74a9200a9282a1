"""Step-4 hand-off file at device speed: same bytes after gunzip as grid_b200.io /
the reference (normalize_mosdepth.py:502-554, find_neighbors.py:81-124), but the N x Rw cells are formatted and
parsed by CUDA kernels (csrc/textio.cu) and gzip runs as independent members on a thread pool (a concatenation of
gzip members is a gzip file; `gzip.open`, which the reference reads with, handles it).

The host keeps what is small: ids, scales, the two header lines, line boundaries (memchr).
"""

from __future__ import annotations

import gzip
import zlib
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import io as gio

MEMBER_BYTES = 8 << 20   # uncompressed text per gzip member (rows are never split)
BATCH_BYTES = 256 << 20  # device/pinned text buffer per formatting or parsing batch


GZIP_LEVEL = 3           # hand-off files are intermediates: favour speed (any level inflates to the same bytes)
_EXTRA_ID = b"GB"        # gzip extra subfield: u32 = bytes of this whole member (header + deflate + trailer)
_HDR_LEN = 10 + 2 + 4 + 4  # fixed header, XLEN, subfield header, subfield payload


def _deflate_member(parts, level: int | None = None) -> bytes:
    """One self-contained gzip member whose header says how long the member is (the idea of BGZF), so that a reader
    can hop from member to member and inflate them in parallel.  Any gzip reader just skips the extra field."""
    co = zlib.compressobj(GZIP_LEVEL if level is None else level, zlib.DEFLATED, -15)  # raw deflate
    crc, size, out = 0, 0, []
    for p in parts:
        crc = zlib.crc32(p, crc)
        size += len(p)
        out.append(co.compress(p))
    out.append(co.flush())
    body = b"".join(out)
    total = _HDR_LEN + len(body) + 8
    head = b"\x1f\x8b\x08\x04" + b"\x00\x00\x00\x00" + b"\x00\xff" + (8).to_bytes(2, "little") \
        + _EXTRA_ID + (4).to_bytes(2, "little") + total.to_bytes(4, "little")
    return head + body + (crc & 0xFFFFFFFF).to_bytes(4, "little") + (size & 0xFFFFFFFF).to_bytes(4, "little")


def _split_members(raw: bytes):
    """[(start, end)] of the members of a file written by _deflate_member, or None if any member lacks the size field."""
    spans, pos, n = [], 0, len(raw)
    while pos < n:
        if n - pos < _HDR_LEN or raw[pos:pos + 4] != b"\x1f\x8b\x08\x04" or raw[pos + 12:pos + 14] != _EXTRA_ID \
                or raw[pos + 10:pos + 12] != b"\x08\x00":
            return None
        total = int.from_bytes(raw[pos + 16:pos + 20], "little")
        if total < _HDR_LEN + 8 or pos + total > n:
            return None
        spans.append((pos, pos + total))
        pos += total
    return spans


def _inflate_member(raw, span) -> bytes:
    a, b = span
    data = zlib.decompress(raw[a + _HDR_LEN:b - 8], -15)
    if zlib.crc32(data) != int.from_bytes(raw[b - 8:b - 4], "little"):
        raise OSError("CRC check failed in a gzip member")
    return data


class _MemberStream:
    """File-like view (readline / read) of a multi-member file whose members are inflated `threads` at a time."""

    def __init__(self, raw: bytes, spans, threads: int):
        self.raw, self.spans, self.next = raw, spans, 0
        self.pool = ThreadPoolExecutor(max_workers=max(1, int(threads)))
        self.ahead = []
        self.buf = b""
        self.depth = 2 * max(1, int(threads))

    def _fill(self) -> bool:
        while self.next < len(self.spans) and len(self.ahead) < self.depth:
            self.ahead.append(self.pool.submit(_inflate_member, self.raw, self.spans[self.next]))
            self.next += 1
        if not self.ahead:
            return False
        self.buf += self.ahead.pop(0).result()
        return True

    def readline(self) -> bytes:
        while True:
            nl = self.buf.find(b"\n")
            if nl >= 0:
                line, self.buf = self.buf[:nl + 1], self.buf[nl + 1:]
                return line
            if not self._fill():
                line, self.buf = self.buf, b""
                return line

    def read(self, size: int) -> bytes:
        parts, got = [self.buf], len(self.buf)
        self.buf = b""
        while got < size and (self.ahead or self.next < len(self.spans)):
            self._fill()
            parts.append(self.buf)
            got += len(self.buf)
            self.buf = b""
        return b"".join(parts)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.pool.shutdown(wait=False, cancel_futures=True)


def _open_text(path, threads: int):
    """Parallel member reader for our own files, plain streaming gzip for anybody else's (the reference's)."""
    with open(path, "rb") as f:
        probe = f.read(_HDR_LEN)
        if len(probe) == _HDR_LEN and probe[:4] == b"\x1f\x8b\x08\x04" and probe[12:14] == _EXTRA_ID:
            raw = probe + f.read()
            spans = _split_members(raw)
            if spans is not None:
                return _MemberStream(raw, spans, threads)
    return gzip.open(path, "rb")


def _fixed3_line(values) -> str:
    return "\t".join(["NA" if v != v else "%.3f" % v for v in np.asarray(values, dtype=np.float64).tolist()])


def write_normalized_output(path, ids, zq, rw: int, mu_sel: np.ndarray, ratio_sel: np.ndarray, raw_means: np.ndarray,
                            threads: int = 8, level: int | None = None) -> None:
    """Step-4 file from the DEVICE matrix zq[N, ldq] (int16 codes, torch CUDA tensor), N = len(ids) rows."""
    import torch

    from ._lib import call, load, ptr

    n = len(ids)
    assert zq.is_cuda and zq.dtype == torch.int16 and zq.shape[0] >= n
    st = torch.cuda.current_stream().cuda_stream
    head = (f"{n}\t{rw}\t" + _fixed3_line(mu_sel) + "\n" + f"{n}\t{rw}\t" + _fixed3_line(ratio_sel) + "\n").encode()
    pool = ThreadPoolExecutor(max_workers=max(1, int(threads)))
    with open(path, "wb") as out:
        if rw == 0 or n == 0:  # nothing for the device to do: rows are "id\tscale\t\n" (join of no fields)
            body = "".join(f"{sid}\t{raw_means[i]:.2f}\t\n" for i, sid in enumerate(ids)).encode()
            out.write(_deflate_member([head, body], level))
            return
        per = int(max(1, min(65535, BATCH_BYTES // (8 * rw))))
        ws = torch.empty(load().gridb200_format_z_ws_bytes(per, rw), dtype=torch.uint8, device=zq.device)
        row_off = torch.empty(per + 1, dtype=torch.int64, device=zq.device)
        text = torch.empty(per * rw * 8, dtype=torch.uint8, device=zq.device)
        host = torch.empty(per * rw * 8, dtype=torch.uint8, pin_memory=True)
        pending = [head]
        pending_bytes = len(head)
        jobs = []
        for r0 in range(0, n, per):
            cnt = min(per, n - r0)
            base = zq[r0:r0 + cnt]
            call("gridb200_format_z_measure", ptr(base), cnt, rw, zq.stride(0), ptr(row_off), ptr(ws), ws.numel(), st)
            call("gridb200_format_z_write", ptr(base), cnt, rw, zq.stride(0), ptr(row_off), ptr(ws), ptr(text),
                 text.numel(), st)
            off = row_off[: cnt + 1].cpu().numpy()  # synchronises the stream
            total = int(off[-1])
            host[:total].copy_(text[:total])
            blob = host[:total].numpy().tobytes()  # detach from the pinned buffer: workers outlive this batch
            view = memoryview(blob)
            for i in range(cnt):
                pre = f"{ids[r0 + i]}\t{raw_means[r0 + i]:.2f}\t".encode()
                pending += [pre, view[off[i]:off[i + 1]]]
                pending_bytes += len(pre) + int(off[i + 1] - off[i])
                if pending_bytes >= MEMBER_BYTES:
                    jobs.append(pool.submit(_deflate_member, pending, level))
                    pending, pending_bytes = [], 0
            while len(jobs) > 4 * max(1, int(threads)):  # bound the text held in flight
                out.write(jobs.pop(0).result())
        if pending:
            jobs.append(pool.submit(_deflate_member, pending, level))
        for j in jobs:
            out.write(j.result())
    pool.shutdown()


class _Fallback(Exception):
    pass


def read_normalized_data(path, device="cuda", threads: int = 8):
    """Step-4 file -> (ids, ratios[Rw] float64, zq[N, ldq] int16 DEVICE tensor, scales[N] float64).
    Files outside the writer's grammar are re-read by grid_b200.io.read_normalized_data (Python float() semantics)."""
    import torch

    try:
        return _read_device(path, torch.device(device), threads)
    except _Fallback:
        ids, ratios, zq, scales = gio.read_normalized_data(path)
        ld = (zq.shape[1] + 7) // 8 * 8
        t = torch.full((len(ids), max(ld, 8)), gio.ZQ_NA, dtype=torch.int16, device=device)
        if zq.size:
            t[:, : zq.shape[1]] = torch.from_numpy(zq).to(device)
        return ids, ratios, t, scales


def _read_device(path, dev, threads):
    import torch

    from ._lib import call, load, ptr

    st = torch.cuda.current_stream().cuda_stream
    ids, scales = [], []
    with _open_text(path, threads) as fh:
        h1 = fh.readline()
        h2 = fh.readline()
        try:
            n_hdr, rw = int(h1.split(b"\t", 2)[0]), int(h1.split(b"\t", 2)[1])
        except (ValueError, IndexError):
            raise _Fallback()
        toks = h2.strip().split(b"\t")[2:]
        ratios = np.array([float("nan") if t in (b"NA", b"nan") else float(t) for t in toks])
        if len(ratios) != rw or rw == 0 or n_hdr <= 0:
            raise _Fallback()
        ld = (rw + 7) // 8 * 8
        zq = torch.full((n_hdr, ld), gio.ZQ_NA, dtype=torch.int16, device=dev)
        stats = torch.zeros(4, dtype=torch.int64, device=dev)
        bad = torch.zeros(4, dtype=torch.int64, device=dev)
        carry = b""
        row = 0
        while True:
            block = fh.read(BATCH_BYTES)
            buf = carry + block
            if not buf:
                break
            end = len(buf) if not block else buf.rfind(b"\n") + 1
            if end == 0:  # no complete line yet
                carry = buf
                continue
            carry = buf[end:]
            offs = [0]
            pos = 0
            while pos < end:
                nl = buf.find(b"\n", pos, end)
                nxt = end if nl < 0 else nl + 1
                t1 = buf.find(b"\t", pos, nxt)
                t2 = buf.find(b"\t", t1 + 1, nxt) if t1 >= 0 else -1
                if t2 < 0:
                    raise _Fallback()
                try:
                    scales.append(float(buf[t1 + 1:t2]))
                except ValueError:
                    raise _Fallback()
                ids.append(buf[pos:t1].decode())
                offs.append(nxt)
                pos = nxt
            cnt = len(offs) - 1
            if row + cnt > n_hdr:
                raise _Fallback()
            t_text = torch.frombuffer(bytearray(buf[:end]), dtype=torch.uint8).to(dev)
            t_off = torch.tensor(offs, dtype=torch.int64, device=dev)
            max_line = int(np.diff(np.asarray(offs)).max())
            for r0 in range(0, cnt, 65535):
                c = min(65535, cnt - r0)
                ws = torch.empty(load().gridb200_parse_z_ws_bytes(c, max_line), dtype=torch.uint8, device=dev)
                call("gridb200_parse_z_rows", ptr(t_text), t_off.data_ptr() + 8 * r0, c, max_line, 2, rw,
                     zq.data_ptr() + (row + r0) * ld * 2, ld, ptr(stats), ptr(ws), ws.numel(), st)
                bad += stats
            row += cnt
            if not block:
                break
        b = bad.cpu().numpy()
        if b[1] or b[2] or b[0] != row * rw:
            raise _Fallback()
    return ids, ratios, zq[:row], np.array(scales)


# ---- step-5 file (save_neighbors, find_neighbors.py:231-267) --------------------------------------------------------
def save_neighbors(path, ids, scales, nbr_idx, nbr_d2, r_use: int, threads: int = 8, level: int | None = None) -> None:
    """The neighbours file from the DEVICE k-best lists (nbr_idx int32 [N, k], nbr_d2 int64 [N, k] exact squared
    distances in 1e-4 units): same bytes after gunzip as grid_b200.io.save_neighbors.  The host contributes one
    "id\\tscale" token per sample; every entry is formatted by gridb200_format_nbrs_*; rows are deflated as independent
    gzip members on a thread pool."""
    import torch

    from ._lib import call, load, ptr

    n = len(ids)
    assert nbr_idx.is_cuda and nbr_idx.dtype == torch.int32 and nbr_d2.dtype == torch.int64 and nbr_idx.shape == nbr_d2.shape
    assert nbr_idx.shape[0] == n and nbr_idx.stride(1) in (0, 1) and nbr_idx.stride(0) == nbr_d2.stride(0)
    dev = nbr_idx.device
    st = torch.cuda.current_stream().cuda_stream
    k = int(nbr_idx.shape[1])
    with open(path, "wb") as out:
        if n == 0:
            out.write(_deflate_member([b""], level))
            return
        toks = [f"{sid}\t{v:.2f}".encode() for sid, v in zip(ids, np.asarray(scales, dtype=np.float64).tolist())]
        tok_off = np.concatenate([[0], np.cumsum([len(t) for t in toks])]).astype(np.int64)
        t_tok = torch.frombuffer(bytearray(b"".join(toks)), dtype=torch.uint8).to(dev)
        t_off = torch.from_numpy(tok_off).to(dev)
        longest = int(np.diff(tok_off).max())
        row_cap = longest + 1 + k * (2 + longest + 24)  # an entry: two tabs, a token, at most 24 bytes of distance
        per = int(max(1, min(65535, n, BATCH_BYTES // row_cap)))
        ws = torch.empty(load().gridb200_format_nbrs_ws_bytes(per), dtype=torch.uint8, device=dev)
        row_off = torch.empty(per + 1, dtype=torch.int64, device=dev)
        text = torch.empty(per * row_cap, dtype=torch.uint8, device=dev)
        host = torch.empty(per * row_cap, dtype=torch.uint8, pin_memory=True)
        pool = ThreadPoolExecutor(max_workers=max(1, int(threads)))
        jobs = []
        ld = nbr_idx.stride(0) if k else 0
        for r0 in range(0, n, per):
            cnt = min(per, n - r0)
            pi = nbr_idx.data_ptr() + 4 * r0 * ld if k else None
            pd = nbr_d2.data_ptr() + 8 * r0 * ld if k else None
            call("gridb200_format_nbrs_measure", ptr(t_tok), ptr(t_off), n, pi, pd, cnt, k, ld, r0, int(r_use), ptr(row_off),
                 ptr(ws), ws.numel(), st)
            call("gridb200_format_nbrs_write", ptr(t_tok), ptr(t_off), n, pi, pd, cnt, k, ld, r0, int(r_use), ptr(row_off),
                 ptr(text), text.numel(), st)
            off = row_off[: cnt + 1].cpu().numpy()  # synchronises the stream
            total = int(off[-1])
            host[:total].copy_(text[:total])
            blob = host[:total].numpy().tobytes()  # detach from the pinned buffer: workers outlive this batch
            view = memoryview(blob)
            a = 0
            while a < cnt:  # members of >= MEMBER_BYTES whole rows
                b = int(np.searchsorted(off, off[a] + MEMBER_BYTES, side="left"))
                b = min(max(b, a + 1), cnt)
                jobs.append(pool.submit(_deflate_member, [view[off[a]:off[b]]], level))
                a = b
            while len(jobs) > 4 * max(1, int(threads)):
                out.write(jobs.pop(0).result())
        for j in jobs:
            out.write(j.result())
        pool.shutdown()


def load_neighbors_device(path, device="cuda", threads: int = 8):
    """Neighbours file -> (ids, scales[N] float64, idx[N, kmax] int32 DEVICE tensor, -1 padded), for files whose every
    "nbr_id\\tnbr_scale" is the "id\\tscale" head of one of the lines (what step 5 writes) and whose ids are distinct.
    Raises _Fallback for anything else; the caller then uses grid_b200.io.load_neighbors (the reference's rules)."""
    import torch

    from ._lib import call, load, ptr

    dev = torch.device(device)
    st = torch.cuda.current_stream().cuda_stream
    ids, scales, heads, batches = [], [], [], []
    kmax = 0
    with _open_text(path, threads) as fh:
        carry = b""
        while True:
            block = fh.read(BATCH_BYTES)
            buf = carry + block
            if not buf:
                break
            if b"\r" in buf:
                raise _Fallback()
            end = len(buf) if not block else buf.rfind(b"\n") + 1
            if end == 0:
                carry = buf
                continue
            carry = buf[end:]
            if not buf[:end].endswith(b"\n"):  # last line of the file without a newline
                buf, end = buf[:end] + b"\n", end + 1
            offs, pos = [0], 0
            while pos < end:
                nxt = buf.find(b"\n", pos, end) + 1
                t1 = buf.find(b"\t", pos, nxt)
                t2 = buf.find(b"\t", t1 + 1, nxt) if t1 >= 0 else -1
                stop = t2 if t2 >= 0 else nxt - 1
                if t1 < 0 or t1 == pos:
                    raise _Fallback()
                try:
                    scales.append(float(buf[t1 + 1:stop]))
                except ValueError:
                    raise _Fallback()
                if buf[t1 + 1:stop].strip() != buf[t1 + 1:stop] or buf[pos:t1].strip() != buf[pos:t1]:
                    raise _Fallback()  # blanks the reference's strip() / float() would remove
                ids.append(buf[pos:t1].decode())
                heads.append(buf[pos:stop])
                fields = buf.count(b"\t", pos, nxt) + 1
                if (fields - 2) % 3:
                    raise _Fallback()
                kmax = max(kmax, (fields - 2) // 3)
                offs.append(nxt)
                pos = nxt
            batches.append((torch.frombuffer(bytearray(buf[:end]), dtype=torch.uint8).to(dev),
                            torch.tensor(offs, dtype=torch.int64, device=dev), int(np.diff(np.asarray(offs)).max())))
            if not block:
                break
    n = len(ids)
    if n == 0 or len(set(ids)) != n:
        raise _Fallback()
    idx = torch.full((n, max(kmax, 1)), -1, dtype=torch.int32, device=dev)
    if kmax == 0:
        return ids, np.array(scales), idx
    tok_off = np.concatenate([[0], np.cumsum([len(h) for h in heads])]).astype(np.int64)
    t_tok = torch.frombuffer(bytearray(b"".join(heads)), dtype=torch.uint8).to(dev)
    t_off = torch.from_numpy(tok_off).to(dev)
    hashes = torch.empty(n, dtype=torch.int64, device=dev)
    call("gridb200_hash_tokens", ptr(t_tok), ptr(t_off), n, ptr(hashes), st)
    hs, order = torch.sort(hashes)
    if n > 1 and bool((hs[1:] == hs[:-1]).any()):
        raise _Fallback()
    hrow = order.to(torch.int32)
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    total = torch.zeros(4, dtype=torch.int64, device=dev)
    row = 0
    for t_text, t_lines, max_line in batches:
        cnt = t_lines.numel() - 1
        for r0 in range(0, cnt, 65535):
            c = min(65535, cnt - r0)
            ws = torch.empty(load().gridb200_parse_z_ws_bytes(c, max_line), dtype=torch.uint8, device=dev)
            n_ent = torch.empty(c, dtype=torch.int32, device=dev)
            call("gridb200_parse_nbrs_rows", ptr(t_text), t_lines.data_ptr() + 8 * r0, c, max_line, ptr(t_tok), ptr(t_off),
                 ptr(hs), ptr(hrow), n, kmax, idx.data_ptr() + 4 * (row + r0) * idx.stride(0), idx.stride(0), ptr(n_ent),
                 ptr(stats), ptr(ws), ws.numel(), st)
            total += stats
        row += cnt
    b = total.cpu().numpy()
    if b[1] or b[2]:
        raise _Fallback()
    return ids, np.array(scales), idx
