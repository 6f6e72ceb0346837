"""Pins the mode bench.py times: NCNET_CONV_MODE=tc16 (fp16 tensor-core operands, fp32 accumulation, fp16 hidden
activations) with SIGNED default-initialised weights, so every ReLU mask bites in forward and backward.

Two references, two questions (oracle/tc16_emulation.py explains the split):

  * ``oracle.tc16_emulation`` in float64 rounds the same tensors to fp16 at the same places as the kernels.  What is
    left is summation order (fp32 in TMEM vs float64) -- plus its one side effect: a stored activation that sits on an
    fp16 rounding boundary can land one fp16 ulp apart (measured: ~1e-4 of the output's max-abs).  A wrong mask stride,
    tap, halo or channel would be an O(1) error.  Bars: forward 3e-4, weight / bias gradients 1e-3.
  * ``oracle.head`` in fp32/fp64 is the reference's own arithmetic.  With fp16 operands (10-bit mantissa, relative
    rounding <= 4.9e-4 per operand, two operands per product, two to three layers) the forward bar on uniform random
    volumes is 1e-3; on real, heavy-tailed volumes (a few large terms dominate a sum, so the roundings do not average
    out) it is 2.5e-3 on the Conv4d stack and 4e-3 after the final MutualMatching (y = x^3/(max_A max_B) up to triples
    the relative error of x).  The fp32 CUDA-core mode is the one that meets the north star's 1e-3 (it is at 1e-5).
  * Gradients through ReLUs: an fp16 rounding moves a pre-activation that sits within ~5e-4 of zero across it, and
    that unit's whole gradient path switches.  With a coherent (positive) output gradient the weight-gradient sums are
    coherent too and one flipped unit is 1/N of a sum: bar 1e-2 against the reference, 1e-3 against the fp16 model.
    With a random-sign output gradient every sum is itself a random walk and one flipped unit is ~1/sqrt(N) of it
    (2e-2 .. 6e-2 on these small volumes); that case is printed with the flip count and bounded at 1e-1.

All metrics are ``nerr`` = max-abs error / max-abs reference.
"""
import pytest
import torch

from conftest import make_head, nerr, record
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
    ([5, 5, 5], [16, 16, 1], (3, 4, 25, 25), 2),     # PF plane: six position chunks -> three CTA pairs per plane (walk16), walk along iA (walk1)
    ([5, 5, 5], [16, 16, 1], (6, 5, 9, 10), 2),      # one chunk per plane: the pair's second CTA has nothing to load
    ([5, 5, 5], [16, 16, 1], (2, 80, 9, 10), 2),     # long jA axis: several work items per persistent CTA (the TMEM ring carries over)
    ([5, 5, 5], [16, 16, 1], (27, 3, 11, 12), 1),    # long iA axis: jA segments in walk16, more outputs than TMEM ring slots in walk1
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
    yo = onc64(x64)
    g_pos = torch.rand(ye.shape, dtype=torch.float64) + 0.5         # coherent output gradient
    g_rnd = torch.randn(ye.shape, dtype=torch.float64)             # random-sign output gradient
    plist = [p for pair in zip(ws64, bs64) for p in pair]

    xg = x.to(cuda).requires_grad_(True)
    ws, bs = nc._params()
    yb = F_.neigh_consensus_branches(xg, ws, bs, True, _tc())
    y = yb[:, :1] + yb[:, 1:]
    flips, total = E.count_relu_flips(onc64, x64.detach())
    res = {}
    for tag, go in (("pos", g_pos), ("rnd", g_rnd)):
        ge = torch.autograd.grad(ye, [x64] + plist, go, retain_graph=True)
        gr = torch.autograd.grad(yo, [x64] + plist, go, retain_graph=True)
        gg = torch.autograd.grad(y, [xg] + list(nc.parameters()), go.float().to(cuda), retain_graph=True)
        res[tag] = ([nerr(a, b) for a, b in zip(gg, ge)], [nerr(a, b) for a, b in zip(gg, gr)])
    print(f"\nstack {ks}/{chs} {dims} B={bsz}: fwd vs fp16 model {nerr(y, ye):.2e}, vs reference {nerr(y, yo):.2e}; "
          f"coherent g: grads vs model {max(res['pos'][0][1:]):.2e} (gx {res['pos'][0][0]:.2e}), vs reference {max(res['pos'][1]):.2e}; "
          f"random g: vs model {max(res['rnd'][0]):.2e}, vs reference {max(res['rnd'][1]):.2e}; "
          f"ReLU flips fp16-vs-exact {flips}/{total}")
    record(f"stack_k{ks[0]}x{len(ks)}_{'x'.join(map(str, dims))}_B{bsz}", {
        "fwd_nerr_vs_fp16_model": nerr(y, ye), "fwd_nerr_vs_reference": nerr(y, yo),
        "grad_nerr_coherent_g_vs_fp16_model": res["pos"][0], "grad_nerr_coherent_g_vs_reference": res["pos"][1],
        "grad_nerr_random_g_vs_fp16_model": res["rnd"][0], "grad_nerr_random_g_vs_reference": res["rnd"][1],
        "grad_order": "x, then (weight, bias) per layer", "relu_flips_fp16_vs_exact": flips, "hidden_units": total})
    assert nerr(y, ye) < 3e-4
    assert nerr(y, yo) < 1e-3
    for e in res["pos"][0][1:]:         # weight / bias gradients, coherent g, against the fp16 model
        assert e < 1e-3
    for e in res["pos"][1][1:]:         # ... and against the reference
        assert e < 1e-2
    # the first-layer INPUT gradient under the coherent g is a heavily cancelling sum over signed weights (and runs
    # on the fp32 CUDA-core kernel with unrounded weights): it is bounded with the random g, like everything else
    for e in res["rnd"][0] + res["rnd"][1]:
        assert e < 1e-1


def _features(b, c, fs, seed):
    from bench import synthetic_pairs            # planted correspondences: the data bench.py times
    return synthetic_pairs(b, c, fs, seed)


def _agreement(m, mo):
    return ((m[5].cpu() == mo[5]) & (m[6].cpu() == mo[6]) & (m[7].cpu() == mo[7]) & (m[8].cpu() == mo[8])).float().mean().item()


def test_tc_head_pf_shape_c1024(F_, cuda, monkeypatch):
    """Whole head at the PF-Pascal shape (25^4, C = 1024) in the benchmarked mode against the fp32 oracle: Conv4d-stack
    output within 2.5e-3, final volume within 4e-3 (module docstring), arg-max matches agree in both directions."""
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
    record("head_pf_25x4_C1024_tc16", {"stack_nerr_vs_reference": e_stack, "head_nerr_vs_reference": e_head,
                                       "head_nerr_vs_fp16_model": e_model, "argmax_agreement_B_to_A": agree[0],
                                       "argmax_agreement_A_to_B": agree[1], "matches_per_direction": 625})
    assert e_stack < 2.5e-3 and e_head < 4e-3 and e_model < 2.5e-3
    assert min(agree) > 0.99


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
        record(f"train5_pf_B2_C1024_{mode}", {"losses": losses, "oracle_losses": ref_losses, "max_abs_dloss": dl,
                                              "step1_grad_nerr_vs_reference": eg})
        assert all(v == v and abs(v) < 1 for v in losses)
        if mode == "fp32":
            assert dl < 2e-3 * scale and max(eg) < 2e-3
        else:
            em = [nerr(a, b) for a, b in zip(grads, model_g)]
            print(f"tc16 step-1 gradient vs fp16 model: {['%.1e' % e for e in em]}; "
                  f"loss vs model {abs(losses[0] - float(lm)):.2e}")
            record("train5_pf_B2_C1024_tc16_vs_fp16_model", {"step1_grad_nerr": em, "step1_loss_abs_diff": abs(losses[0] - float(lm))})
            assert abs(losses[0] - float(lm)) < 2e-3 * scale
            assert max(em) < 2e-2
            assert dl < 5e-2 * scale and max(eg) < 5e-2
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
    record("inloc_1600px_fp16_head", {"nerr_vs_fp32_reference": e, "match_agreement_B_to_A": agree[0],
                                      "match_agreement_A_to_B": agree[1]})
    # every tensor of this head is fp16 (the reference's half_precision mode): bar 1e-2 against the fp32 oracle
    assert e < 1e-2 and min(agree) > 0.9
