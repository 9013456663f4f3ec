"""Makes the reference's real tskit Python package importable for the CPU interop tests, where it
can be had: its C extension from oracle/_ref/tskit_py (`make -C oracle tskit`, part of build()),
its pure-Python part from the reference tree itself.  MSPB_TSKIT_PATH overrides.  Returns None where
neither is there (the GPU box)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_EXT = os.path.join(ROOT, "oracle", "_ref", "tskit_py")
_PKG = "/root/reference/git-submodules/tskit/python"


def load():
    if not os.environ.get("MSPB_TSKIT_PATH") and os.path.isdir(os.path.join(_PKG, "tskit")) \
            and not os.path.exists(os.path.join(_EXT, ".stamp")):
        # the reference tree is here but its extension has not been compiled yet (build() does this too)
        import subprocess

        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "tskit"], check=False)
    if os.environ.get("MSPB_TSKIT_PATH"):
        extra = [os.environ["MSPB_TSKIT_PATH"]]
    elif os.path.isdir(_EXT) and os.path.isdir(os.path.join(_PKG, "tskit")):
        extra = [_EXT, _PKG]
    else:
        extra = []
    for p in extra:
        if p not in sys.path:
            sys.path.append(p)
    saved = sys.dont_write_bytecode
    sys.dont_write_bytecode = True  # the reference tree is read-only: no __pycache__ attempts in it
    try:
        return importlib.import_module("tskit")
    except ImportError:
        return None
    finally:
        sys.dont_write_bytecode = saved
