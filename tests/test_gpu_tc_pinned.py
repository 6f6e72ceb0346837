"""Pins the mode bench.py times: NCNET_CONV_MODE=tc16 (fp16 tensor-core operands, fp32 accumulation, fp16 hidden
activations) with SIGNED default-initialised weights, so every ReLU mask bites in forward and backward.

Two references, two questions (oracle/tc16_emulation.py explains the split):

  * ``oracle.tc16_emulation`` in float64 rounds the same tensors to fp16 at the same places as the kernels.  What is
    left is summation order (fp32 in TMEM vs float64), so a wrong mask stride, tap, halo or channel would show as an
    O(1) error.  Bars: forward 1e-4, weight / bias gradients 1e-3.
  * ``oracle.head`` in fp32/fp64 is the reference's own arithmetic.  Bars: forward of the Conv4d stack 1e-3 (the north
    star's bound for the fp32-accumulate mode), 2e-3 after the final MutualMatching (y = x^3/(max_A max_B) triples the
    relative error of x).  Gradients through ReLUs cannot be held to 1e-3: an fp16 rounding moves a pre-activation that
    sits within ~5e-4 of zero across it, and that unit's whole gradient path switches.  The tests count those flips
    and bound the gradient error by 5e-2 for a random (incoherent) output gradient.

All metrics are ``nerr`` = max-abs error / max-abs reference.
"""
import pytest
import torch

from conftest import make_head, nerr
from oracle import head as O
from oracle import tc16_emulation as E

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def F_(cuda):
    from ncnet_b200 import functional
    return functional


def _tc():
    from ncnet_b200._cabi import CONV_TC16
    return CONV_TC16


def _double(onc, ks, chs):
    o64 = O.NeighConsensus(ks, chs).double()
    o64.load_state_dict({k: v.double() for k, v in onc.state_dict().items()})
    return o64


@pytest.mark.parametrize("ks,chs,dims,bsz", [
    ([5, 5, 5], [16, 16, 1], (3, 4, 25, 25), 2),     # PF plane: two chunks per plane -> CTA pairs; CB16 mask staged in smem
    ([5, 5, 5], [16, 16, 1], (6, 5, 9, 10), 2),      # one chunk per plane, two CTAs per SM
    ([3, 3], [16, 1], (5, 6, 20, 13), 2),            # InLoc stack
    ([3, 3], [16, 1], (3, 4, 50, 37), 1),            # InLoc-pooled (1600 px) plane
])
def test_tc_signed_stack_gradients(F_, cuda, ks, chs, dims, bsz):
    import ncnet_b200 as nb
    torch.manual_seed(41)
    onc = O.NeighConsensus(ks, chs)                  # signed default init: about half of the hidden units are off
    x = torch.rand(bsz, 1, *dims)
    nc = nb.NeighConsensus(True, ks, chs)
    nc.load_state_dict(onc.state_dict())

    onc64 = _double(onc, ks, chs)
    ws64, bs64 = E._params(onc64)
    x64 = x.double().requires_grad_(True)
    ye = E.neigh_consensus(x64, ws64, bs64, True)
    go = torch.randn(ye.shape, dtype=torch.float64)
    plist = [p for pair in zip(ws64, bs64) for p in pair]
    ge = torch.autograd.grad(ye, [x64] + plist, go)
    yo = onc64(x64)
    gr = torch.autograd.grad(yo, [x64] + plist, go)

    xg = x.to(cuda).requires_grad_(True)
    ws, bs = nc._params()
    yb = F_.neigh_consensus_branches(xg, ws, bs, True, _tc())
    y = yb[:, :1] + yb[:, 1:]
    gg = torch.autograd.grad(y, [xg] + list(nc.parameters()), go.float().to(cuda))

    flips, total = E.count_relu_flips(onc64, x64.detach())
    e_model = [nerr(a, b) for a, b in zip(gg, ge)]
    e_ref = [nerr(a, b) for a, b in zip(gg, gr)]
    print(f"\nstack {ks}/{chs} {dims} B={bsz}: fwd vs fp16 model {nerr(y, ye):.2e}, vs reference {nerr(y, yo):.2e}; "
          f"grads vs model {max(e_model[1:]):.2e} (gx {e_model[0]:.2e}), vs reference {max(e_ref):.2e}; "
          f"ReLU flips fp16-vs-exact {flips}/{total}")
    assert nerr(y, ye) < 1e-4
    assert nerr(y, yo) < 1e-3
    assert e_model[0] < 3e-3            # first-layer input gradient: CUDA-core kernel with unrounded weights
    for e in e_model[1:]:
        assert e < 1e-3
    for e in e_ref:
        assert e < 5e-2


def _features(b, c, fs, seed):
    g = torch.Generator().manual_seed(seed)
    one = lambda: O.feature_l2norm(torch.relu(torch.randn(b, c, fs, fs, generator=g)))
    return one(), one()


def _agreement(m, mo):
    return ((m[5].cpu() == mo[5]) & (m[6].cpu() == mo[6]) & (m[7].cpu() == mo[7]) & (m[8].cpu() == mo[8])).float().mean().item()


def test_tc_head_pf_shape_c1024(F_, cuda, monkeypatch):
    """Whole head at the PF-Pascal shape (25^4, C = 1024) in the benchmarked mode: Conv4d-stack output within 1e-3 of
    the fp32 oracle, final volume within 2e-3, matches agree; against the fp16 model both are at fp32-rounding level."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(42)
    fa, fb = _features(1, 1024, 25, 43)
    onc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    with torch.no_grad():
        x0 = O.mutual_matching(O.feature_correlation(fa, fb, "4D", False))
        ref_stack = onc(x0)
        ref = O.mutual_matching(ref_stack)
        model = E.head_forward(fa.double(), fb.double(), _double(onc, [5, 5, 5], [16, 16, 1]))
    net = make_head(nb, [5, 5, 5], [16, 16, 1], onc.state_dict())
    with torch.no_grad():
        fa_g, fb_g = fa.to(cuda), fb.to(cuda)
        x0_g = nb.MutualMatching(net.FeatureCorrelation(fa_g, fb_g))
        stack = net.NeighConsensus(x0_g)
        out = net.match_features(fa_g, fb_g)
    e_stack, e_head, e_model = nerr(stack, ref_stack), nerr(out, ref), nerr(out, model)
    agree = []
    for inv in (False, True):
        m = nb.corr_to_matches(out, do_softmax=True, return_indices=True, invert_matching_direction=inv)
        mo = O.corr_to_matches(ref, do_softmax=True, return_indices=True, invert_matching_direction=inv)
        agree.append(_agreement(m, mo))
    print(f"\nTC16 head 25^4 C=1024: stack nerr {e_stack:.2e}, head nerr {e_head:.2e}, vs fp16 model {e_model:.2e}, "
          f"arg-max agreement B->A {agree[0]:.4f} A->B {agree[1]:.4f}")
    assert e_stack < 1e-3 and e_head < 2e-3 and e_model < 1e-4
    assert min(agree) > 0.95


def _train(module, step_fn, steps):
    params = [p for p in module.parameters() if p.requires_grad]
    opt = torch.optim.Adam(params, lr=5e-4)
    losses, first_grads = [], None
    for s in range(steps):
        opt.zero_grad()
        loss = step_fn()
        loss.backward()
        if s == 0:
            first_grads = [p.grad.detach().cpu().clone() for p in params]
        opt.step()
        losses.append(float(loss.detach()))
    return losses, first_grads


def test_tc_training_steps_match_reference(F_, cuda, monkeypatch):
    """Five Adam steps of the weak loss (train.py:110-172) on a PF-shaped batch of 2 pairs, C = 1024, signed default
    weights: fp32 mode and TC16 mode on the GPU against the fp32 oracle, plus the step-1 gradient of TC16 mode against
    the fp16 model, plus bit-reproducibility of the TC16 run."""
    import ncnet_b200 as nb
    from ncnet_b200.train_util import weak_loss_from_features
    steps = 5
    fa, fb = _features(2, 1024, 25, 44)
    torch.manual_seed(45)
    init = O.NeighConsensus([5, 5, 5], [16, 16, 1]).state_dict()

    onc = O.NeighConsensus([5, 5, 5], [16, 16, 1])
    onc.load_state_dict(init)
    ref_losses, ref_g = _train(onc, lambda: O.weak_loss(fa, fb, onc), steps)

    onc64 = O.NeighConsensus([5, 5, 5], [16, 16, 1]).double()
    onc64.load_state_dict({k: v.double() for k, v in init.items()})
    lm = E.weak_loss(fa.double(), fb.double(), onc64)
    model_g = torch.autograd.grad(lm, list(onc64.parameters()))

    fa_g, fb_g = fa.to(cuda), fb.to(cuda)
    got = {}
    for mode in ("fp32", "tc16"):
        monkeypatch.setenv("NCNET_CONV_MODE", mode)
        net = make_head(nb, [5, 5, 5], [16, 16, 1], init)
        got[mode] = _train(net.NeighConsensus, lambda: weak_loss_from_features(net, fa_g, fb_g), steps)
    scale = max(abs(v) for v in ref_losses)
    for mode, (losses, grads) in got.items():
        dl = max(abs(a - b) for a, b in zip(losses, ref_losses))
        eg = [nerr(a, b) for a, b in zip(grads, ref_g)]
        print(f"\n{mode}: losses {['%.4e' % v for v in losses]} (oracle {['%.4e' % v for v in ref_losses]}), "
              f"max |dloss| {dl:.2e}, step-1 gradient nerr per tensor {['%.1e' % e for e in eg]}")
        assert all(v == v and abs(v) < 1 for v in losses)
        if mode == "fp32":
            assert dl < 1e-3 * scale + 1e-8 and max(eg) < 2e-3
        else:
            em = [nerr(a, b) for a, b in zip(grads, model_g)]
            print(f"tc16 step-1 gradient vs fp16 model: {['%.1e' % e for e in em]}; "
                  f"loss vs model {abs(losses[0] - float(lm)):.2e}")
            assert abs(losses[0] - float(lm)) < 1e-3 * scale + 1e-8
            assert max(em) < 2e-3
            assert dl < 5e-2 * scale + 1e-7 and max(eg) < 5e-2
    # the same seed twice gives the same bits: no floating-point atomics anywhere in the step
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    net = make_head(nb, [5, 5, 5], [16, 16, 1], init)
    again = _train(net.NeighConsensus, lambda: weak_loss_from_features(net, fa_g, fb_g), steps)
    assert again[0] == got["tc16"][0]
    for a, b in zip(again[1], got["tc16"][1]):
        assert torch.equal(a, b)


def test_inloc_1600_real_shape(F_, cuda, monkeypatch):
    """BASELINE.json configs[2] at its real size: 100x74 features (1600 px), fp16, maxpool4d k=2, NC 3-3 / 16-1,
    corr_to_matches both directions with relocalisation -- against the fp32 oracle (about 10 s of CPU)."""
    import ncnet_b200 as nb
    monkeypatch.setenv("NCNET_CONV_MODE", "tc16")
    torch.manual_seed(46)
    g = torch.Generator().manual_seed(47)
    fa = O.feature_l2norm(torch.relu(torch.randn(1, 1024, 100, 74, generator=g)))
    fb = O.feature_l2norm(torch.relu(torch.randn(1, 1024, 74, 100, generator=g)))
    onc = O.NeighConsensus([3, 3], [16, 1])
    with torch.no_grad():
        ref, rdelta = O.head_forward(fa.half().float(), fb.half().float(), onc, relocalization_k_size=2)
    net = make_head(nb, [3, 3], [16, 1], onc.state_dict(), reloc=2, half=True)
    with torch.no_grad():
        out, delta = net.match_features(fa.half().to(cuda), fb.half().to(cuda))
    assert out.dtype == torch.float16 and tuple(out.shape) == (1, 1, 50, 37, 37, 50)
    e = nerr(out.float(), ref)
    agree = []
    for inv in (False, True):
        m = nb.corr_to_matches(out, delta4d=delta, k_size=2, do_softmax=True, scale="positive", return_indices=True,
                               invert_matching_direction=inv)
        mo = O.corr_to_matches(ref, delta4d=rdelta, k_size=2, do_softmax=True, scale="positive", return_indices=True,
                               invert_matching_direction=inv)
        agree.append(_agreement(m, mo))
    print(f"\nInLoc 1600 px head (fp16): nerr {e:.2e}, match agreement B->A {agree[0]:.4f}, A->B {agree[1]:.4f}")
    assert e < 5e-3 and min(agree) > 0.9
