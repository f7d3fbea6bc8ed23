"""Recipe for ``oracle/_ref/`` — TEST / BASELINE INFRASTRUCTURE, NOT PRODUCT CODE.

The reference (simeonjunker/retr) is plain Python with no build step and no packaging, so "building" it
for the CPU baseline means placing its hot-path modules where ``bench.py --impl reference`` can import
them on the GPU box (``/root/reference`` does not exist there):

    models/*.py  eval_utils/decode.py  configuration_template.py  LICENSE   ->   oracle/_ref/

``oracle/_ref/`` is git-ignored (the sources never enter this repository's history) but travels to the
GPU box with the snapshot, exactly like a compiled ``.so``.  Run by ``__graft_entry__.build()`` whenever
``/root/reference`` is present; nothing under ``retr_b200/`` imports the result.

    python oracle/make_ref.py [/root/reference]
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = ["configuration_template.py", "LICENSE", "eval_utils/decode.py"]
DIRS = ["models"]


def make(src="/root/reference"):
    if not os.path.isdir(src):
        return None
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(os.path.join(DST, "eval_utils"))
    open(os.path.join(DST, "eval_utils", "__init__.py"), "w").close()
    for f in FILES:
        shutil.copy2(os.path.join(src, f), os.path.join(DST, f))
    for d in DIRS:
        os.makedirs(os.path.join(DST, d), exist_ok=True)
        for f in sorted(os.listdir(os.path.join(src, d))):
            if f.endswith(".py"):
                shutil.copy2(os.path.join(src, d, f), os.path.join(DST, d, f))
    with open(os.path.join(DST, "PROVENANCE"), "w") as f:
        f.write(f"copied unmodified from {src} by oracle/make_ref.py\n")
    return DST


def load():
    """Import the unmodified reference modules from oracle/_ref (None when the directory is absent).
    The only patch is the one the survey documents: ``models.backbone.is_main_process`` -> False so that
    Backbone.__init__ does not try to download ImageNet weights (models/backbone.py:87-88)."""
    if not os.path.isfile(os.path.join(DST, "models", "caption.py")):
        return None
    if DST not in sys.path:
        sys.path.insert(0, DST)
    import importlib
    bb = importlib.import_module("models.backbone")
    bb.is_main_process = lambda: False
    return {"caption": importlib.import_module("models.caption"),
            "utils": importlib.import_module("models.utils"),
            "decode": importlib.import_module("eval_utils.decode"),
            "Config": importlib.import_module("configuration_template").Config}


if __name__ == "__main__":
    print(make(*sys.argv[1:2]))
