"""Stage the UNMODIFIED reference hot-path files under baseline/_ref/ so that they travel to the GPU box.

    python -m oracle.stage_ref

The reference is pure Python and ships no packaging (no setup.py / pyproject.toml), so `pip install
--target baseline/_ref /root/reference` cannot work; this recipe copies the seven files the path consists of,
byte for byte, keeping their relative paths.  baseline/_ref/ is git-ignored (never part of the history, never
product source) but not gpurun-ignored.  TEST / MEASUREMENT INFRASTRUCTURE: only oracle/_ref_shims.py reads
the staged tree (bench.py --impl reference / --impl torch-gpu; tests never do).  Nothing under scflow_b200/
may import it.
"""
import hashlib
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("SCFLOW_REFERENCE_SRC", "/root/reference")
DST = os.path.join(ROOT, "baseline", "_ref")
FILES = (
    "models/decoder/raft_decoder.py", "models/decoder/scflow_decoder.py", "models/utils/corr_lookup.py",
    "models/utils/pose.py", "models/head/pose_head.py", "models/encoder/raft_encoder.py",
    "models/backbone/resnet.py",
)


def stage(verbose: bool = True) -> bool:
    """Copy the files when the reference tree is present; returns whether baseline/_ref is complete."""
    if os.path.isfile(os.path.join(SRC, FILES[1])):
        for f in FILES:
            dst = os.path.join(DST, f)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(os.path.join(SRC, f), dst)
        with open(os.path.join(DST, "MANIFEST.txt"), "w") as m:
            for f in FILES:
                m.write(f"{hashlib.sha256(open(os.path.join(DST, f), 'rb').read()).hexdigest()}  {f}\n")
    ok = all(os.path.isfile(os.path.join(DST, f)) for f in FILES)
    if verbose:
        print(("staged " if ok else "NOT staged: ") + DST)
    return ok


if __name__ == "__main__":
    sys.exit(0 if stage() else 1)
