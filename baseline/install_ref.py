"""Copy the unmodified reference package into baseline/_ref and verify checksums."""

from __future__ import annotations

import hashlib
import pathlib
import shutil
import sys

HERE = pathlib.Path(__file__).resolve().parent
SUMS = {
    "mpitree/__init__.py": "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855",
    "mpitree/tree/__init__.py": "55e86ae279b49fc80c9094e336de3e788509e59aefff07077a27a08a875cfd45",
    "mpitree/tree/_base.py": "fb6478ced3a5bf41f8a83a957f2958caf7a4ebb890936ab3b10401a37f9d20d2",
    "mpitree/tree/decision_tree.py": "def4db0d8e787dbe8e3709ab073b0eac4e2c4aeae5486f1aed7410b15579e4c7",
}


def verify(root: pathlib.Path) -> bool:
    for rel, want in SUMS.items():
        p = root / rel
        if not p.exists() or hashlib.sha256(p.read_bytes()).hexdigest() != want:
            return False
    return True


def main(src: str = "/root/reference") -> int:
    dst = HERE / "_ref"
    if verify(dst):
        print("baseline/_ref already holds the unmodified reference")
        return 0
    srcp = pathlib.Path(src)
    if not (srcp / "mpitree").is_dir():
        print(f"reference source not found at {src}", file=sys.stderr)
        return 1
    dst.mkdir(exist_ok=True)
    shutil.rmtree(dst / "mpitree", ignore_errors=True)
    shutil.copytree(srcp / "mpitree", dst / "mpitree")
    ok = verify(dst)
    print("installed" if ok else "checksum mismatch", dst)
    return 0 if ok else 2


if __name__ == "__main__":
    sys.exit(main(*sys.argv[1:]))
