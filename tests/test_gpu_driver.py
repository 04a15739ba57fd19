"""Batched window driver on the GPU: the two sliding-window loops of the reference
(scripts/test_holisticbody.py:239-262, scripts/demo.py:116-181) over the CUDA plans."""
import pytest
import torch

from probtalk_b200 import synthetic as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pipe():
    from probtalk_b200 import driver as D
    return D.HolisticPipeline(S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5), "cuda")


def test_windows_eval_feedback_and_known_frames(pipe):
    """330-frame clips = two windows [0,180) and [150,330): the second window sees the first one's last 30 frames
    as known context and must reproduce them exactly; frames the caller marks known come back bit-exact."""
    B, t = 3, 330
    aud = S.synthetic_audio(B, t, 91).cuda()
    gt = S.synthetic_poses(B, t, 92).cuda()
    mask = torch.zeros(B, 1, t, device="cuda")
    mask[:, :, 100:130] = 1                              # an inpainting span inside the first window
    labels = torch.tensor([0, 1, 2], dtype=torch.int32, device="cuda")
    pipe.generator = torch.Generator(device="cuda").manual_seed(7)
    out = pipe.windows_eval(aud, gt, mask, labels)
    assert out.shape == (B, 376, t) and bool(torch.isfinite(out).all())
    assert torch.equal(out[:, :, 100:130], gt[:, :, 100:130])
    # replay window by window: the overlap of window 2 equals window 1's output (refiner returns gt*mask there)
    pipe.generator = torch.Generator(device="cuda").manual_seed(7)
    w1, _ = pipe.window(aud[..., :180].contiguous(), gt[..., :180].contiguous(), mask[..., :180].contiguous(), labels)
    w1 = w1.clone()
    assert torch.equal(out[:, :, :100], w1[:, :, :100])
    assert torch.equal(out[:, :, 150:180], w1[:, :, 150:180])
    # the generated part is not constant and differs between samples
    assert out[:, :, 200:].std() > 0.05 and (out[0] - out[1]).abs().max() > 0.1


def test_windows_demo_long_form(pipe):
    """demo.py generalised: 600 frames -> 4 predictor windows (180, stride 150) and 5 refiner windows (176, stride 146)."""
    from probtalk_b200 import driver as D
    B, t = 2, 600
    aud = S.synthetic_audio(B, t, 93).cuda()
    labels = torch.tensor([3, 0], dtype=torch.int32, device="cuda")
    pipe.generator = torch.Generator(device="cuda").manual_seed(9)
    out, codes = pipe.windows_demo(aud, labels)
    assert out.shape == (B, 376, t) and bool(torch.isfinite(out).all())
    wins = D.demo_window_bounds(t, 180)
    steps = [((e - s_) // 2 // 2 // 2) for s_, e in wins]          # 180 -> 22 token steps, 150 -> 18
    assert len(wins) == 4 and steps == [22, 22, 22, 18]
    assert codes.shape == (B, steps[0] + sum(n - 3 for n in steps[1:]), 4)   # demo.py:151 drops 3 steps per later window
    assert int(codes.min()) >= 0 and int(codes.max()) < 128
    # deterministic for a fixed noise generator
    pipe.generator = torch.Generator(device="cuda").manual_seed(9)
    out2, codes2 = pipe.windows_demo(aud, labels)
    assert torch.equal(codes, codes2) and torch.equal(out, out2)


def test_graph_replay_equals_eager(pipe):
    """CUDA-graph replay of a window gives bit-identical results to eager launches with the same noise."""
    from probtalk_b200 import engine as E
    B, T = 2, 180
    aud, gt = S.synthetic_audio(B, T, 95).cuda(), S.synthetic_poses(B, T, 96).cuda()
    mask = torch.zeros(B, 1, T, device="cuda")
    mask[:, :, :30] = 1
    labels = torch.tensor([1, 2], dtype=torch.int32, device="cuda")
    g = torch.Generator().manual_seed(3)
    u = torch.rand(8, B, 88, generator=g).cuda()
    gum = -torch.log(-torch.log(torch.rand(8, B, 88, generator=g).clamp_(1e-10, 1 - 1e-7))).cuda()
    outs = []
    for use_graph in (False, True):
        pl = E.PredictorPlan(pipe.PW, pipe.VW, B, T, use_graph=use_graph)
        rp = E.RefinerPlan(pipe.RW, B, T, T, use_graph=use_graph)
        for _ in range(3):                                # call 1 eager warm-up, call 2 captures, call 3 replays
            poses, codes = pl.run(aud, gt, mask, labels, noise=(u, gum))
            ref = rp.run(aud, gt, mask, poses, labels)
        outs.append((poses.clone(), codes.clone(), ref.clone()))
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)


def test_second_device_in_one_process():
    """Per-device kernel set-up (dynamic shared-memory opt-ins, SM count): a process that runs plans on cuda:0 and then on
    cuda:1 (the reference wraps its models in nn.DataParallel) must get the same results on both."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from probtalk_b200 import driver as D
    B, T = 2, 180
    aud, gt = S.synthetic_audio(B, T, 95), S.synthetic_poses(B, T, 96)
    mask = torch.zeros(B, 1, T)
    mask[:, :, :30] = 1
    labels = torch.tensor([1, 2], dtype=torch.int32)
    g = torch.Generator().manual_seed(3)
    u = torch.rand(8, B, 88, generator=g)
    gum = -torch.log(-torch.log(torch.rand(8, B, 88, generator=g).clamp_(1e-10, 1 - 1e-7)))
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        pipe = D.HolisticPipeline(S.make_vq_state(0), S.make_predictor_state(1), S.make_refiner_state(5), dev)
        out, codes = pipe.window(aud.to(dev), gt.to(dev), mask.to(dev), labels.to(dev), noise=(u.to(dev), gum.to(dev)))
        x = S.synthetic_poses(1, 300, 7).to(dev)
        from probtalk_b200 import engine as E
        enc_codes = E.EncoderPlan(pipe.VW, 1, 300).run(x).cpu()
        outs.append((out.cpu(), codes.cpu(), enc_codes))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
