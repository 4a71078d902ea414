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


def test_gru_and_flow_head_are_cuda_graph_capturable():
    """K5 / K5b: two pack launches + eight GEMM launches (+ two for the flow head), tensor maps encoded on the host at
    call time, nothing synchronises: a captured unit replays bit-identically on new inputs in its static buffers."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import torch.nn as nn
    import ppmstereo_b200 as P
    dev = torch.device("cuda")

    def c(ci, co, k):
        return nn.Conv3d(ci, co, k, padding=tuple(x // 2 for x in k))

    class Gru(nn.Module):                                   # the unit's layer shapes (include/ppm_b200.h, K5)
        def __init__(self):
            super().__init__()
            self.convz1 = nn.Sequential(c(512, 128, (1, 1, 15)), nn.GELU(), c(128, 128, (1, 1, 5)))
            self.convr1 = nn.Sequential(c(512, 128, (1, 1, 15)), nn.GELU(), c(128, 128, (1, 1, 5)))
            self.convq1 = c(512, 128, (1, 1, 5))
            self.convz2, self.convr2, self.convq2 = (c(512, 128, (1, 5, 1)) for _ in range(3))
            self.convz3, self.convr3, self.convq3 = (c(512, 128, (5, 1, 1)) for _ in range(3))

    class Head(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv1, self.conv2 = c(128, 256, (3, 3, 3)), c(256, 2, (3, 3, 3))

    torch.manual_seed(4)
    gru, head = P.SepConvGRU3D(Gru().to(dev)), P.FlowHead3D(Head().to(dev))
    g = torch.Generator(device=dev).manual_seed(5)
    h = torch.tanh(torch.randn(1, 128, 4, 10, 24, generator=g, device=dev))
    x = torch.randn(1, 384, 4, 10, 24, generator=g, device=dev)

    def unit():
        net = gru(h, x)
        return net, head(net)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        unit()                                              # warm-up: function attributes, weight packing, workspaces
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = unit()
    for seed in (6, 7):
        gg = torch.Generator(device=dev).manual_seed(seed)
        h.copy_(torch.tanh(torch.randn(h.shape, generator=gg, device=dev)))
        x.copy_(torch.randn(x.shape, generator=gg, device=dev))
        graph.replay()
        torch.cuda.synchronize()
        got = [o.clone() for o in outs]
        want = unit()
        torch.cuda.synchronize()
        for a, b in zip(got, want):
            assert torch.equal(a, b)
