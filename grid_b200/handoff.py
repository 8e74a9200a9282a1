"""In-process, device-resident hand-off between steps 4 -> 5 -> 6 ("optional binary side-car +
device-resident hand-off keyed by output path").

`grid wgs` runs the steps one after the other in ONE process (pipeline.py:66-93), each step reading the text file
the previous one wrote.  The files are still written -- byte-compatible, so any step can be run by either
implementation, now or later -- but when the next step finds that the file it is about to parse is exactly the one
this process has just written (same real path, size, mtime and inode), it takes the arrays that are still sitting
on the GPU instead of inflating and parsing them again.  Anything else (file touched, replaced, written by another
process or by the reference) falls back to reading the file.  Disable with config key gpu.device_handoff: false.
"""

from __future__ import annotations

import os

_CACHE: dict[str, tuple[tuple, dict]] = {}


def _signature(path) -> tuple:
    st = os.stat(path)
    return (st.st_size, st.st_mtime_ns, st.st_ino)


def put(path, **payload) -> None:
    """Remember what was just written to `path` (call after the file is closed).  One entry per path."""
    _CACHE[os.path.realpath(path)] = (_signature(path), payload)


def take(path) -> dict | None:
    """The payload for `path` if the file on disk is still the one `put` saw; the entry is consumed either way."""
    entry = _CACHE.pop(os.path.realpath(path), None)
    if entry is None:
        return None
    try:
        if _signature(path) != entry[0]:
            return None
    except OSError:
        return None
    return entry[1]


def clear() -> None:
    _CACHE.clear()


def enabled(config) -> bool:
    return bool((config.get("gpu") or {}).get("device_handoff", True))
