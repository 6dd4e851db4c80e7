"""Recipe (test / benchmark infrastructure, NOT a product path): stage the UNMODIFIED reference sampler so that it can
travel to the GPU box and be timed there as the CPU baseline and as the stock-PyTorch-on-CUDA baseline.

The reference is pure Python with no build step (no setup.py / pyproject), so "building" it is copying the handful of
files its sampling path imports, byte for byte, from /root/reference/diffusion into oracle/_ref/ (git-ignored: the
sources never enter this repository's history; not gpurun-ignored: they travel like a built .so).  Run in the build
container (done by __graft_entry__.build() whenever /root/reference exists):

    python oracle/make_ref.py

Files staged (paths under /root/reference/diffusion/): diffuser/models/{__init__,diffusion,temporal,helpers,
sampling_config}.py -- GaussianDiffusion.p_sample_loop, TemporalUnet.forward and everything they call -- the
diffuser/utils modules that `import diffuser.utils` pulls in, and metrics/fid_score.py (MotionFID).  A manifest with
sha256 sums is written next to them so a reader can check that nothing was edited.
"""
import hashlib
import json
import os
import shutil
import sys

SRC = "/root/reference/diffusion"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = [
    "diffuser/__init__.py",
    "diffuser/models/__init__.py", "diffuser/models/diffusion.py", "diffuser/models/temporal.py",
    "diffuser/models/helpers.py", "diffuser/models/sampling_config.py",
    "diffuser/utils/__init__.py", "diffuser/utils/serialization.py", "diffuser/utils/training.py",
    "diffuser/utils/progress.py", "diffuser/utils/config.py", "diffuser/utils/arrays.py", "diffuser/utils/logger.py",
    "diffuser/utils/timer.py",
    "metrics/fid_score.py",
]


def stage(verbose=True):
    if not os.path.isdir(SRC):
        if verbose:
            print(f"{SRC} not present: nothing staged (oracle/_ref keeps whatever an earlier run put there)")
        return False
    manifest = {}
    for rel in FILES:
        src = os.path.join(SRC, rel)
        if not os.path.exists(src):
            if rel.endswith("__init__.py"):      # namespace package in the reference: nothing to copy
                continue
            raise FileNotFoundError(src)
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    json.dump({"source": SRC, "files": manifest}, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1)
    if verbose:
        print(f"staged {len(manifest)} reference files into {DST}")
    return True


def available():
    return os.path.exists(os.path.join(DST, "diffuser", "models", "diffusion.py"))


def import_reference():
    """Imports the staged reference (oracle/_ref) and returns (GaussianDiffusion, TemporalUnet).  Only bench.py's
    reference / baseline legs call this."""
    import contextlib
    import io
    if not available():
        raise RuntimeError("oracle/_ref is empty: run `python oracle/make_ref.py` in the build container")
    if DST not in sys.path:
        sys.path.insert(0, DST)
    sys.dont_write_bytecode = True
    with contextlib.redirect_stdout(io.StringIO()):
        from diffuser.models.diffusion import GaussianDiffusion
        from diffuser.models.temporal import TemporalUnet
    return GaussianDiffusion, TemporalUnet


if __name__ == "__main__":
    stage()
