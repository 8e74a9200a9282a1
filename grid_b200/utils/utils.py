"""The few helpers of the reference's grid/utils/utils.py the three steps use (utils.py:13-43, :76-111)."""
from contextlib import contextmanager
from pathlib import Path


def log(console, msg, style=None):
    if console:
        console.print(msg, style=style) if style else console.print(msg)
    else:
        print(msg)


@contextmanager
def progress_bar(console=None, total=1, description="Working"):
    """Same call shape as the reference's rich progress bar; plain (no themed styles needed)."""

    class _Bar:
        def update(self, *a, **k):
            pass

        def advance(self, *a, **k):
            pass

    yield _Bar(), 0


def get_samples(samples_file):
    with open(samples_file) as fh:
        return [ln.strip() for ln in fh if ln.strip()]


def setup_output_file(output_file, chrom, start, end) -> Path:
    path = Path(output_file).expanduser()
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text(f"Sample\t{chrom}:{start}-{end}\n")
    return path
