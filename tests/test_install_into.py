"""The drop-in boundary driven by the REFERENCE's own loader: after ``install_into(nets)`` the
unmodified ``nets/ai1.py::init_model`` (getattr(nets, name)(args, config) -> torch.load -> load_state_dict) must build
the probtalk_b200 wrappers from a checkpoint that the reference's wrappers saved.  CPU only: no kernel runs."""
import os
import types

import pytest
import torch

from oracle import ref_harness as RH
from probtalk_b200 import synthetic as S

pytestmark = pytest.mark.skipif(not RH.reference_available(), reason="reference tree not present")


def _same(a, b, path=""):
    assert set(a.keys()) == set(b.keys()), (path, set(a.keys()) ^ set(b.keys()))
    for k in a:
        if torch.is_tensor(a[k]):
            assert a[k].dtype == b[k].dtype and torch.equal(a[k].cpu(), b[k].cpu()), path + k


def test_reference_init_model_builds_drop_in_wrappers(tmp_path):
    import probtalk_b200.nets as PN
    st = RH.load_reference()
    nets = st["nets"]
    import nets.ai1 as ref_ai1                      # the reference module, unmodified
    vq, pred, ref = S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5)
    # checkpoints written by the REFERENCE wrappers' own state_dict() (Trainer.py:201-216 layout)
    ckpts = {}
    for name, w in (("s2g_body_vqt", RH.build_vqt(vq)), ("s2g_body_predictor", RH.build_predictor(vq, pred)),
                    ("s2g_body_refiner", RH.build_refiner(vq, ref))):
        sd = w.state_dict()
        sd["generator_optim"] = None               # the optimizer state is irrelevant to inference
        path = str(tmp_path / (name + ".pth"))
        torch.save({"generator": sd, "epoch": 0, "global_steps": 0}, path)
        ckpts[name] = (path, sd)
    previous = PN.install_into(nets)
    try:
        cfg = RH._config("transformer.json")
        cfg.Model.vq_path = ckpts["s2g_body_vqt"][0]
        args = types.SimpleNamespace(gpu=0, infer=True)
        for name in ("s2g_body_predictor", "s2g_body_refiner"):
            m = ref_ai1.init_model(name, ckpts[name][0], args, cfg)
            assert type(m) is getattr(PN, name)
            mine, theirs = m.state_dict(), ckpts[name][1]
            for part in mine:
                if part == "generator_optim":
                    continue
                _same(mine[part], {k.replace("module.", ""): v for k, v in theirs[part].items()}, part + ".")
            # the PQ-VAE was fetched from config.Model.vq_path in the constructor (predictor.py:81-83)
            _same(m.VQ.state_dict(), {k: v for k, v in vq.items()}, "VQ.")
        vcfg = RH._config("vq.json")
        m = ref_ai1.init_model("s2g_body_vqt", ckpts["s2g_body_vqt"][0], args, vcfg)
        assert type(m) is PN.s2g_body_vqt
        _same(m.state_dict()["VQ"], ckpts["s2g_body_vqt"][1]["VQ"], "VQ.")
        # All_In_One_Model of the reference chains the drop-in wrappers the same way
        aio = ref_ai1.All_In_One_Model(None, None, ["s2g_body_predictor", "s2g_body_refiner"],
                                       [ckpts["s2g_body_predictor"][0], ckpts["s2g_body_refiner"][0]], "cuda", args, cfg)
        assert [type(x) for x in aio.body_model] == [PN.s2g_body_predictor, PN.s2g_body_refiner]
    finally:
        PN.uninstall(nets, previous)
    assert nets.s2g_body_predictor is previous["s2g_body_predictor"]
