"""TEST INFRASTRUCTURE -- loads the unmodified reference extension built by oracle/build_ref.py (oracle/_ref) under
an alias module name.  Used by tests/util.py, __graft_entry__.smoke() and bench.py's reference arm only."""
import importlib.util
import os
import sys

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "diff_surfel_rasterization")


def load_ref():
    """The reference's python wrapper + _C.so as module `ref_dsr`, or None when oracle/_ref is not built."""
    if "ref_dsr" in sys.modules:
        return sys.modules["ref_dsr"]
    init = os.path.join(REF_DIR, "__init__.py")
    if not (os.path.exists(init) and os.path.exists(os.path.join(REF_DIR, "_C.so"))):
        return None
    spec = importlib.util.spec_from_file_location("ref_dsr", init, submodule_search_locations=[REF_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_dsr"] = mod
    spec.loader.exec_module(mod)
    return mod
