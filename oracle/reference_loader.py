"""Import and build the UNMODIFIED reference model (FasterProcess/sam2_opt) in this container.

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/make_golden.py``, by the CPU tests that pin the
oracle against the real reference, by the GPU tests that drive the reference's own predictors
through ``speedup("b200")`` and by ``bench.py``'s reference / baseline legs.  The sources are
``/root/reference`` in the build container and the byte-identical copy under ``oracle/_ref``
(made by ``oracle/build_ref.py``, git-ignored) on the GPU box.  No product module imports this.

The reference needs ``hydra``/``omegaconf``/``iopath`` only for its builders
(reference ``sam2/sam2/build_sam.py:11-13``, ``sam2/sam2/__init__.py:7-11``,
``modeling/backbones/hieradet.py:14``); they are not installed here, so tiny stubs are
registered and the model is instantiated from the YAML
(``sam2/sam2/configs/sam2.1/sam2.1_hiera_l.yaml``) by a recursive ``_target_`` walker
with the overrides of ``build_sam.py:110-131``.
"""
import importlib
import os
import re
import sys
import types

_CFG = "sam2/sam2/configs/sam2.1/sam2.1_hiera_l.yaml"
_VENDORED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _pick_root() -> str:
    """/root/reference in the build container; on the GPU box the byte-identical copy that `oracle/build_ref.py` made
    under the git-ignored `oracle/_ref/` (it travels with the working tree)."""
    env = os.environ.get("SAM2_REFERENCE_ROOT")
    for cand in ([env] if env else []) + ["/root/reference", _VENDORED]:
        if cand and os.path.isfile(os.path.join(cand, _CFG)):
            return cand
    return env or "/root/reference"


REF_ROOT = _pick_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, _CFG))


def kind() -> str:
    """'reference' = the unmodified reference sources (in place or the oracle/_ref copy), 'absent' otherwise."""
    return "reference" if available() else "absent"


def _install_stubs():
    if "hydra" in sys.modules and getattr(sys.modules["hydra"], "_sam2b_stub", False):
        return
    hydra = types.ModuleType("hydra")
    hydra._sam2b_stub = True
    hydra.initialize_config_module = lambda *a, **k: None
    hydra.compose = lambda *a, **k: None
    hcore = types.ModuleType("hydra.core")
    gh = types.ModuleType("hydra.core.global_hydra")

    class GlobalHydra:
        @staticmethod
        def instance():
            return types.SimpleNamespace(is_initialized=lambda: True)

    gh.GlobalHydra = GlobalHydra
    hutils = types.ModuleType("hydra.utils")
    hutils.instantiate = lambda *a, **k: None
    om = types.ModuleType("omegaconf")
    om.OmegaConf = object
    io = types.ModuleType("iopath")
    ioc = types.ModuleType("iopath.common")
    iof = types.ModuleType("iopath.common.file_io")
    iof.g_pathmgr = None
    sys.modules.update({
        "hydra": hydra, "hydra.core": hcore, "hydra.core.global_hydra": gh,
        "hydra.utils": hutils, "omegaconf": om,
        "iopath": io, "iopath.common": ioc, "iopath.common.file_io": iof,
    })


def _coerce(v):
    # PyYAML reads "1e-6" as str; OmegaConf reads it as float.
    if isinstance(v, str) and re.fullmatch(r"[-+]?\d+(\.\d*)?[eE][-+]?\d+", v):
        return float(v)
    return v


def _instantiate(node):
    if isinstance(node, dict):
        kw = {k: _instantiate(v) for k, v in node.items() if k != "_target_"}
        if "_target_" in node:
            mod, _, name = node["_target_"].rpartition(".")
            return getattr(importlib.import_module(mod), name)(**kw)
        return kw
    if isinstance(node, list):
        return [_instantiate(v) for v in node]
    return _coerce(node)


def build_reference(kind="video", seed=0, fill_hole_area=0):
    """kind: 'video' -> SAM2VideoPredictor, 'base' -> SAM2Base (for SAM2ImagePredictor)."""
    import torch
    import yaml

    assert available(), f"reference not found under {REF_ROOT}"
    _install_stubs()
    p = os.path.join(REF_ROOT, "sam2")
    if p not in sys.path:
        sys.path.insert(0, p)
    os.environ.setdefault("SAM2_VERSION_TRACK", "official")
    with open(os.path.join(REF_ROOT, _CFG)) as f:
        cfg = yaml.safe_load(f)["model"]
    if kind == "video":
        cfg["_target_"] = "sam2.sam2_video_predictor.SAM2VideoPredictor"
        # build_sam.py:119-130 (build_sam2_video_predictor hydra overrides)
        cfg["sam_mask_decoder_extra_args"] = dict(
            dynamic_multimask_via_stability=True,
            dynamic_multimask_stability_delta=0.05,
            dynamic_multimask_stability_thresh=0.98,
        )
        cfg["binarize_mask_from_pts_for_mem_enc"] = True
        cfg["fill_hole_area"] = fill_hole_area  # 8 upstream; needs sam2._C (absent) -> 0
    else:
        # build_sam.py:81-88 (build_sam2 overrides)
        cfg["sam_mask_decoder_extra_args"] = dict(
            dynamic_multimask_via_stability=True,
            dynamic_multimask_stability_delta=0.05,
            dynamic_multimask_stability_thresh=0.98,
        )
    torch.manual_seed(seed)
    model = _instantiate(cfg)
    return model.eval()
