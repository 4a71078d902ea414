"""SURVEY 8(b) 'Threading / streams': every entry launches on the caller's stream, allocates nothing and never
synchronises, so one update iteration of the hot path (lookup + pick + pack + aggregate + upsample) can be captured
in a CUDA graph and replayed on new inputs.  Replay must equal the eager result bit for bit."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_iteration_is_cuda_graph_capturable():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ppmstereo_b200 as P
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(9)
    T, h, w, c = 8, 24, 40, 128

    def rn(*s):
        return torch.randn(*s, generator=g, device=dev)

    f1, f2 = rn(T, 256, h, w).bfloat16(), rn(T, 256, h, w).bfloat16()
    blk = P.CorrBlock1D(f1, f2)
    mem = P.PPMMemory(rn(1, c, T, h, w), rn(1, c, T, h, w), None, top_k=5)
    sim = mem.similarity()
    beta = torch.tensor([0.5], device=dev)
    # static inputs of one iteration
    flow, value, mf = torch.zeros(T, 2, h, w, device=dev), rn(1, c, T, h, w), rn(1, c, T, h, w)
    unc, up_mask = torch.rand(T, 1, h, w, generator=g, device=dev), rn(T, 144, h, w)
    strive = torch.ones(1, 1, T, T, device=dev)

    def iteration():
        corr = blk(flow)
        mfg, mask = mem.read(value, mf, sim, strive, unc, beta)
        return corr, mfg, mask, P.convex_upsample(flow, up_mask)

    def new_inputs(seed):
        gg = torch.Generator(device=dev).manual_seed(seed)
        flow[:, 0] = torch.rand(T, h, w, generator=gg, device=dev) * -30.0
        value.copy_(torch.randn(value.shape, generator=gg, device=dev))
        mf.copy_(torch.randn(mf.shape, generator=gg, device=dev))
        unc.copy_(torch.rand(unc.shape, generator=gg, device=dev))

    new_inputs(1)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        iteration()                                    # warm-up outside capture (function attributes, tensor maps)
    torch.cuda.current_stream().wait_stream(side)
    strive.fill_(1.0)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = iteration()
    for seed in (2, 3, 4):                              # strive_time carries over between replays, as in the loop
        new_inputs(seed)
        before = strive.clone()
        graph.replay()
        torch.cuda.synchronize()
        got = [o.clone() for o in outs]
        after = strive.clone()
        strive.copy_(before)
        want = iteration()
        torch.cuda.synchronize()
        assert torch.equal(strive, after)
        for a, b in zip(got, want):
            assert torch.equal(a, b)
