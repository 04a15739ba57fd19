"""Run the UNMODIFIED ProbTalk reference on CPU (test infrastructure only).

This module is TEST INFRASTRUCTURE.  It exists to pin ``oracle/probtalk_oracle.py``
against the real reference and to generate the committed golden vectors in
``tests/golden/``, and to run the reference arm of ``bench.py`` (``--impl reference`` / ``cpu_baseline``).  It imports
``/root/reference`` (read-only, present only in the authoring container) or, on the GPU box, the unmodified snapshot
``oracle/_ref/reference`` made by ``oracle/build_ref.py``; nothing on the product path, in the ``-m gpu`` tests or in
``smoke()`` may import it.

The reference hard-codes CUDA devices, needs a handful of absent third-party modules
at import time and breaks on torch >= 2.0; the shims below (none of which edit the
reference tree) are the ones listed in SURVEY.md section 8(c) / Appendix B:

  1. cwd + sys.path = /root/reference (relative opens at import time)
  2. stub modules: matplotlib(.pyplot), librosa, textgrid, interval, smplx,
     python_speech_features, torchaudio.sox_effects
  3. 'cuda' device strings -> cpu (TorchFunctionMode) + no-op torch.cuda.synchronize
  4. DeepSupervisionLayer.forward must swallow tgt_is_causal / memory_is_causal
  5. fake data_root with 4 speaker dirs; a vq checkpoint for config.Model.vq_path
"""
from __future__ import annotations

import contextlib
import importlib.machinery
import os
import sys
import tempfile
import types

import torch

_SNAPSHOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "reference")     # oracle/build_ref.py
REFERENCE_ROOT = os.environ.get("PROBTALK_REFERENCE", "/root/reference")
if not os.path.isdir(os.path.join(REFERENCE_ROOT, "nets")) and os.path.isdir(os.path.join(_SNAPSHOT, "nets")):
    REFERENCE_ROOT = _SNAPSHOT            # the GPU box: the unmodified snapshot that travelled with the repository


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "nets"))


class _CudaToCpu(torch.overrides.TorchFunctionMode):
    """Rewrite every 'cuda*' device argument to 'cpu' (shim 3)."""

    @staticmethod
    def _fix(v):
        if isinstance(v, str) and v.startswith("cuda"):
            return "cpu"
        if isinstance(v, torch.device) and v.type == "cuda":
            return torch.device("cpu")
        return v

    def __torch_function__(self, func, types_, args=(), kwargs=None):
        kwargs = kwargs or {}
        if getattr(func, "__name__", "") == "cuda":          # tensor.cuda()
            return args[0]
        args = tuple(self._fix(a) for a in args)
        kwargs = {k: self._fix(v) for k, v in kwargs.items()}
        return func(*args, **kwargs)


_STATE = {}


def load_reference():
    """Import the reference's ``nets`` package under the shims; returns a namespace."""
    if "nets" in _STATE:
        return _STATE
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    import transformers  # noqa: F401  (must be imported before librosa is stubbed)

    for name in ["matplotlib", "matplotlib.pyplot", "librosa", "textgrid", "interval", "smplx",
                 "python_speech_features", "torchaudio.sox_effects"]:
        if name in sys.modules:
            continue
        m = types.ModuleType(name)
        m.__spec__ = importlib.machinery.ModuleSpec(name, None)
        sys.modules[name] = m
    sys.modules["interval"].Interval = object
    sys.modules["torchaudio.sox_effects"].apply_effects_tensor = None
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]

    torch.cuda.synchronize = lambda *a, **k: None
    old_cwd = os.getcwd()
    os.chdir(REFERENCE_ROOT)
    sys.path.insert(0, REFERENCE_ROOT)
    mode = _CudaToCpu()
    mode.__enter__()                       # stays active for the life of the process
    try:
        import nets  # the reference package
        from nets.inpainting import PTransformer as PT
        from trainer.config import load_JsonConfig
    finally:
        os.chdir(old_cwd)

    orig_forward = PT.DeepSupervisionLayer.forward

    def forward(self, *a, tgt_is_causal=None, memory_is_causal=None, **k):
        return orig_forward(self, *a, **k)

    PT.DeepSupervisionLayer.forward = forward

    tmp = tempfile.mkdtemp(prefix="probtalk_ref_")
    data_root = os.path.join(tmp, "data_root")
    for spk in ["oliver", "chemistry", "seth", "conan"]:
        os.makedirs(os.path.join(data_root, spk))
    _STATE.update(nets=nets, PT=PT, load_JsonConfig=load_JsonConfig, tmp=tmp, data_root=data_root,
                  mode=mode)
    return _STATE


def _config(name):
    st = load_reference()
    cfg = st["load_JsonConfig"](os.path.join(REFERENCE_ROOT, "config", name))
    cfg.Data.data_root = st["data_root"]
    return cfg


@contextlib.contextmanager
def _quiet():
    with open(os.devnull, "w") as dn, contextlib.redirect_stdout(dn):
        yield


def build_vqt(vq_state=None):
    """reference ``s2g_body_vqt`` (nets/inpainting/vq_teacher.py:23) on CPU."""
    st = load_reference()
    cfg = _config("vq.json")
    args = types.SimpleNamespace(gpu="cpu", infer=True)
    with _quiet():
        w = st["nets"].s2g_body_vqt(args, cfg)
    if vq_state is not None:
        w.VQ.load_state_dict(vq_state)
    w.VQ.eval()
    return w


def _vq_ckpt(vq_state):
    st = load_reference()
    if vq_state is None:
        vq_state = build_vqt().VQ.state_dict()
    path = os.path.join(st["tmp"], "vq_ckpt.pth")
    torch.save({"generator": {"VQ": vq_state}}, path)
    return path


def build_predictor(vq_state=None, state=None):
    """reference ``s2g_body_predictor`` (nets/inpainting/predictor.py:27) on CPU.
    ``state`` = {'AudEnc':..., 'Predictor':..., 'TextEnc':...} (optional)."""
    st = load_reference()
    cfg = _config("transformer.json")
    cfg.Model.vq_path = _vq_ckpt(vq_state)
    args = types.SimpleNamespace(gpu="cpu", infer=True)
    with _quiet():
        w = st["nets"].s2g_body_predictor(args, cfg)
    if state is not None:
        w.load_state_dict({k: v for k, v in state.items() if k != "generator_optim"})
    for m in (w.AudEnc, w.TextEnc, w.Predictor, w.VQ):
        m.eval()
    return w


def build_refiner(vq_state=None, state=None):
    """reference ``s2g_body_refiner`` (nets/inpainting/refiner.py:32) on CPU."""
    st = load_reference()
    cfg = _config("transformer.json")
    cfg.Model.vq_path = _vq_ckpt(vq_state)
    args = types.SimpleNamespace(gpu="cpu", infer=True)
    with _quiet():
        w = st["nets"].s2g_body_refiner(args, cfg)
    if state is not None:
        w.load_state_dict({k: v for k, v in state.items() if k != "generator_optim"})
    w.RTransformer.eval()
    w.VQ.eval()
    return w
