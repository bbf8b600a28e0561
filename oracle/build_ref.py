"""Recipe for oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot): the reference's own
`app/file_parser.py`, taken unmodified from where it lies under /root/reference, plus the one stub
module it needs to import here (`docx`, never used on the .csv / .txt path).

    python oracle/build_ref.py        # no-op when /root/reference is not mounted

TEST / BASELINE INFRASTRUCTURE ONLY: `bench.py --impl reference` and the tests load it through
oracle/ref_runner.py to time and to cross-check the reference itself; the product never imports it.
Nothing under oracle/_ref/ is committed.
"""
from __future__ import annotations

import shutil
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF = Path("/root/reference")
OUT = HERE / "_ref"


def build() -> bool:
    src = REF / "app" / "file_parser.py"
    if not src.exists():
        return (OUT / "app" / "file_parser.py").exists()
    (OUT / "app").mkdir(parents=True, exist_ok=True)
    (OUT / "stubs").mkdir(parents=True, exist_ok=True)
    shutil.copyfile(src, OUT / "app" / "file_parser.py")
    (OUT / "app" / "__init__.py").write_text("")
    (OUT / "stubs" / "docx.py").write_text('"""empty stand-in: app/file_parser.py imports docx at module level"""\n')
    return True


if __name__ == "__main__":
    print("oracle/_ref ready" if build() else "reference not mounted and no oracle/_ref yet")
