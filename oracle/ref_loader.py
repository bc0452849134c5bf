"""Import the reference's own modules (TEST / BASELINE INFRASTRUCTURE ONLY).

Looks for the unmodified reference files in oracle/_ref/ (staged by oracle/vendor_ref.py; this is what exists on
the GPU box) and then in /root/reference (build container).  `vae-hpo.py` has a hyphen in its name, so it is loaded
with importlib; its `from utils import *` must resolve to the REFERENCE's utils.py, so the directory goes to the
front of sys.path and a `utils` module imported earlier from somewhere else is dropped first."""
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CANDIDATES = (os.path.join(HERE, "_ref"), "/root/reference")


def ref_dir():
    for d in CANDIDATES:
        if os.path.exists(os.path.join(d, "vae-hpo.py")) and os.path.exists(os.path.join(d, "utils.py")):
            return d
    return None


def load_vae_hpo():
    """The reference's vae-hpo.py as a module (VAE, loss_function, train, test, run, setup_ddp ...), or None."""
    d = ref_dir()
    if d is None:
        return None
    if d in sys.path:
        sys.path.remove(d)
    sys.path.insert(0, d)
    m = sys.modules.get("utils")
    if m is not None and os.path.dirname(os.path.abspath(getattr(m, "__file__", ""))) != d:
        del sys.modules["utils"]
    spec = importlib.util.spec_from_file_location("ref_vae_hpo", os.path.join(d, "vae-hpo.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
