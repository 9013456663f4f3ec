"""Imports the REFERENCE's pure-Python ``msprime/intervals.py`` (and the ``core.py`` it uses for its
text table) in this container, where the compiled extension cannot be built: ``_msprime`` is
replaced by an inert stand-in.  Used by tests/golden/make_ratemap_golden.py only; needs
/root/reference; nothing at run time imports this."""
import importlib.util
import sys
import types

REF = "/root/reference/msprime"


def load():
    pkg = types.ModuleType("msprime")
    pkg.__path__ = []
    ext = types.ModuleType("msprime._msprime")
    ext.unset_gsl_error_handler = lambda: None
    pkg._msprime = ext
    stubs = {"msprime": pkg, "msprime._msprime": ext}
    saved = {k: sys.modules.get(k) for k in ("msprime", "msprime._msprime", "msprime.core", "msprime.intervals")}
    sys.modules.update(stubs)
    try:
        mods = {}
        for name in ("core", "intervals"):
            spec = importlib.util.spec_from_file_location(f"msprime.{name}", f"{REF}/{name}.py")
            mod = importlib.util.module_from_spec(spec)
            sys.modules[f"msprime.{name}"] = mod
            setattr(pkg, name, mod)
            spec.loader.exec_module(mod)
            mods[name] = mod
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mods["intervals"]
