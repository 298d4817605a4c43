"""Recipe that makes the UNMODIFIED reference travel to the GPU box.

    python oracle/make_ref.py            # /root/reference/{sailrecon,train} -> oracle/_ref/{sailrecon,train}

TEST INFRASTRUCTURE ONLY.  `oracle/_ref/` is git-ignored (no reference source ever enters the history) but NOT
gpurun-ignored, so the byte-identical copy ships with the snapshot exactly like the built `.so` does.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s reference / parity / library-baseline legs import from it; the product
package never does.  The reference is pure Python (no native code to compile), so the "build" is a verbatim copy of
the `.py` files plus a manifest of their sha256 digests; `verify()` re-checks the digests against /root/reference when
that tree is present.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DST = os.path.join(HERE, "_ref")
SUBTREES = ("sailrecon", "train")


def _sha(path: str) -> str:
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_DST, "sailrecon", "models", "aggregator.py"))


def make(force: bool = False) -> str | None:
    """Copy the reference's Python sources verbatim.  Returns the destination, or None when /root/reference is absent
    (GPU box: the prebuilt copy that travelled with the snapshot is used as is)."""
    if not os.path.isdir(os.path.join(REF_SRC, "sailrecon")):
        return REF_DST if available() else None
    manifest = {}
    for sub in SUBTREES:
        src_root = os.path.join(REF_SRC, sub)
        for d, _, files in os.walk(src_root):
            for fn in files:
                if not fn.endswith(".py"):
                    continue
                src = os.path.join(d, fn)
                rel = os.path.relpath(src, REF_SRC)
                dst = os.path.join(REF_DST, rel)
                digest = _sha(src)
                manifest[rel] = digest
                if not force and os.path.exists(dst) and _sha(dst) == digest:
                    continue
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copyfile(src, dst)
    with open(os.path.join(REF_DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": REF_SRC, "files": manifest}, f, indent=0, sort_keys=True)
    return REF_DST


def verify() -> bool:
    """Every copied file is byte-identical to the manifest (and to /root/reference when present)."""
    with open(os.path.join(REF_DST, "MANIFEST.json")) as f:
        files = json.load(f)["files"]
    for rel, digest in files.items():
        if _sha(os.path.join(REF_DST, rel)) != digest:
            return False
        src = os.path.join(REF_SRC, rel)
        if os.path.exists(src) and _sha(src) != digest:
            return False
    return True


def import_path() -> str:
    """sys.path entry under which `import sailrecon` resolves to the unmodified reference."""
    if not available():
        raise RuntimeError("oracle/_ref is absent: run `python oracle/make_ref.py` in the build container")
    return REF_DST


if __name__ == "__main__":
    dst = make(force="--force" in sys.argv)
    print(dst, "verified" if dst and verify() else "unavailable")
