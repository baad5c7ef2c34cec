"""Recipe for oracle/_ref/: the reference's OWN Python sources for the hot path, copied unmodified from /root/reference at
build time so that `bench.py --impl reference` and the `cpu_baseline` leg can time the reference itself on the GPU box's host
cores (where /root/reference does not exist).  TEST / BASELINE INFRASTRUCTURE: nothing under f5e_tts_b200/ imports it.

oracle/_ref/ is git-ignored (it never enters history) but not gpurun-ignored (it travels with the snapshot, like the built
.so).  Only the packages `from f5_tts.model import CFM, DiT` pulls in are copied; third-party packages that are not installed
(torchdiffeq, x_transformers, ...) stay behind oracle/shims/.

    python oracle/make_ref.py            # no-op with a message when /root/reference is absent
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/f5_tts"
DST = os.path.join(HERE, "_ref", "f5_tts")
PACKAGES = ("model", "durpred", "ppg")  # what `import f5_tts.model` imports (see the module list in DESIGN.md 4)


def make() -> bool:
    if not os.path.isdir(SRC):
        print(f"oracle/make_ref.py: {SRC} not present (GPU box): keeping whatever oracle/_ref holds")
        return os.path.isdir(DST)
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST)
    shutil.copy2(os.path.join(SRC, "__init__.py"), DST) if os.path.exists(os.path.join(SRC, "__init__.py")) else open(
        os.path.join(DST, "__init__.py"), "w").close()
    n = 0
    for pkg in PACKAGES:
        for root, dirs, files in os.walk(os.path.join(SRC, pkg)):
            dirs[:] = [d for d in dirs if d != "__pycache__"]
            out = os.path.join(DST, os.path.relpath(root, SRC))
            os.makedirs(out, exist_ok=True)
            for f in files:
                if f.endswith(".py"):
                    shutil.copy2(os.path.join(root, f), out)
                    n += 1
    with open(os.path.join(HERE, "_ref", "README"), "w") as f:
        f.write("Unmodified copies of /root/reference/src/f5_tts/{model,durpred,ppg}/**/*.py made by oracle/make_ref.py.\n"
                "Not part of the repository (git-ignored); used only as the timed CPU baseline.\n")
    print(f"oracle/make_ref.py: copied {n} reference files into {DST}")
    return True


def import_reference():
    """-> (CFM, DiT) classes of the unmodified reference from oracle/_ref (None, None when it has not been made)."""
    if not os.path.isdir(DST):
        return None, None
    for p in (os.path.join(HERE, "_ref"), os.path.join(HERE, "shims")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from f5_tts.model import CFM, DiT
    return CFM, DiT


if __name__ == "__main__":
    make()
