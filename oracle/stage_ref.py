"""Stage the UNMODIFIED reference modules of the hot path where the GPU box can import them.  TEST / BENCH INFRASTRUCTURE ONLY.

/root/reference exists only in the build container.  This recipe copies the reference's own files for the path -
avion/models/{transformer,model_clip,utils,model_videomae}.py and avion/losses/losses.py (plus the empty package markers) -
byte for byte into oracle/_ref/avion/ (git-ignored: no reference source enters the history; NOT gpurun-ignored: it travels
with the snapshot like the built .so files).  __graft_entry__.build() runs it whenever /root/reference is present.
oracle/ref_import.py then imports the reference from /root/reference if it is there, else from oracle/_ref, with the same
two shims (timm.models.layers: 3-4 symbols; clip: empty module).

Users: bench.py --impl reference / cpu_baseline (the reference's CLIP + ClipLoss on the host cores, kind "reference") and
tools/compare_same_box.py (the reference's flash-attn block on the same B200).  The product never imports it.
"""
import filecmp
import os
import shutil

SRC = "/root/reference/avion"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "avion")
FILES = ["__init__.py", "models/__init__.py", "models/transformer.py", "models/model_clip.py", "models/utils.py",
         "models/model_videomae.py", "losses/__init__.py", "losses/losses.py"]


def stage(verbose=False):
    """Copy (only what changed).  Returns the staged root or None when the reference tree is absent."""
    if not os.path.isdir(SRC):
        return None
    for rel in FILES:
        src, dst = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or not filecmp.cmp(src, dst, shallow=False):
            shutil.copyfile(src, dst)
            if verbose:
                print("staged", rel)
    return os.path.dirname(DST)


if __name__ == "__main__":
    print(stage(verbose=True))
