"""oracle/ref_import.py -- TEST INFRASTRUCTURE (golden-vector generation only).

Imports the UNMODIFIED reference codec modules straight from /root/reference
(read-only, present only in the build container, never on the GPU box) so that
`oracle/gen_golden.py` can run the real `DMC` / `IntraNoAR` and pin the oracle.

Two shims, neither touching the reference tree:
  * `pytorch_msssim` is not installed; `common_model.py:7` imports `MS_SSIM`
    and constructs it at :29 but never calls it -> a stub module.
  * `entropy_models.py:13,19` import `.MLCodec_rans` / `.MLCodec_CXX` relative
    to the read-only package dir -> the modules built by `oracle/Makefile`
    (reference C++ compiled unmodified into oracle/_ref/) are pre-registered in
    `sys.modules` under those dotted names.
"""
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("VCM_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_SO = os.path.join(_HERE, "_ref")


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "DCVC_HEM", "src", "models"))


def load_ref_native(name):
    """Load oracle/_ref/<name>*.so (MLCodec_rans | MLCodec_CXX) as a module."""
    import glob
    hits = sorted(glob.glob(os.path.join(_REF_SO, name + "*.so")))
    if not hits:
        raise ImportError(f"oracle/_ref/{name}*.so missing - run `make -C oracle ref`")
    spec = importlib.util.spec_from_file_location(name, hits[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def import_reference():
    """Return (DMC, IntraNoAR, stream_helper module) of the real reference."""
    if not reference_available():
        raise ImportError(f"reference tree not found at {REF_ROOT}")
    if "pytorch_msssim" not in sys.modules:
        import torch
        stub = types.ModuleType("pytorch_msssim")

        class MS_SSIM(torch.nn.Module):  # constructed, never called
            def __init__(self, *a, **k):
                super().__init__()

        stub.MS_SSIM = MS_SSIM
        stub.ms_ssim = lambda *a, **k: None
        sys.modules["pytorch_msssim"] = stub
    for name in ("MLCodec_rans", "MLCodec_CXX"):
        dotted = "DCVC_HEM.src.entropy_models." + name
        if dotted not in sys.modules:
            sys.modules[dotted] = load_ref_native(name)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from DCVC_HEM.src.models.video_model import DMC
    from DCVC_HEM.src.models.image_model import IntraNoAR
    from DCVC_HEM.src.utils import stream_helper
    return DMC, IntraNoAR, stream_helper
