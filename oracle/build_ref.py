"""TEST / BASELINE INFRASTRUCTURE -- not part of the product path.

Recipe that vendors the reference's own Python modules for the hot path into `oracle/_ref/` (git-ignored, NOT
gpurun-ignored: it travels to the GPU box like a built .so) so that `bench.py --impl reference` and the `cpu_baseline`
leg time the REFERENCE ITSELF (cpu_baseline.kind = "reference") instead of the oracle port.  The reference is pure
Python: nothing is compiled; the files are copied byte for byte from /root/reference where they lie and are
never committed.  Only models/, options/, util/ and dct/dct_native.py (+ the two pure-Python modules it imports) are needed to build `Pix2PixHDModel`
(models/models.py:4-20 -> models/pix2pixHD_model.py -> models/networks.py, models/mdct.py, util/util.py,
dct/dct_native.py); four absent third-party imports are stubbed at import time by oracle/ref_runner.py.

    python oracle/build_ref.py            # in the build container (also run by __graft_entry__.build())"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("P2P_REFERENCE", "/root/reference")
DST = os.path.join(HERE, "_ref")
WANTED = ["models", "options", "util"] + [os.path.join("dct", f) for f in
                                          ("__init__.py", "dct_native.py", "discrete_spectral_transform.py", "torch_fft_api.py")]


def build(verbose=True):
    """-> True if oracle/_ref is in place (copied now or already there), False when /root/reference is absent"""
    if not os.path.isdir(REF):
        return os.path.isfile(os.path.join(DST, "models", "pix2pixHD_model.py"))
    n = 0
    for item in WANTED:
        src = os.path.join(REF, item)
        if os.path.isdir(src):
            for root, dirs, files in os.walk(src):
                dirs[:] = [d for d in dirs if d not in ("__pycache__", "ParallelWaveGAN")]     # use_hifigan_D is out of scope
                for f in files:
                    if not f.endswith(".py"):
                        continue
                    rel = os.path.relpath(os.path.join(root, f), REF)
                    out = os.path.join(DST, rel)
                    os.makedirs(os.path.dirname(out), exist_ok=True)
                    shutil.copyfile(os.path.join(root, f), out)
                    n += 1
        elif os.path.isfile(src):
            out = os.path.join(DST, item)
            os.makedirs(os.path.dirname(out), exist_ok=True)
            shutil.copyfile(src, out)
            n += 1
    if verbose:
        print(f"oracle/_ref: {n} reference files copied from {REF}")
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
