"""GPU aid: rounding flow of PCBlock4_Deep_nopool_res (models/core/ppmtereo_update.py:91-116) under bf16 autocast on
torch/cuDNN: for the 1x1 convs and the depthwise convs, which candidate reproduces the bf16 result bit for bit?"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn.functional as F
from oracle import ref_import
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
pc = upd.PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(1)
B, H, W = 4, 46, 80
corr = torch.randn(B, 36, H, W, generator=g, device="cuda") * 2
r16 = lambda t: t.to(torch.bfloat16).float()

def variants(conv, inp, groups=1):
    w, b = r16(conv.weight), conv.bias
    acc = F.conv2d(r16(inp).double(), w.double(), None, padding=conv.padding, groups=groups).float()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ref = conv(inp)
    assert ref.dtype == torch.bfloat16
    bb = b.view(1, -1, 1, 1)
    out = {"bf16(acc+bf16 b)": r16(acc + r16(bb)), "bf16(acc+fp32 b)": r16(acc + bb),
           "bf16(bf16(acc)+bf16 b)": r16(r16(acc) + r16(bb))}
    return ref.float(), {k: round(float((v == ref.float()).float().mean()), 5) for k, v in out.items()}

with torch.no_grad():
    print("ffn1[0] 1x1 36->54 ", variants(pc.ffn1[0], corr)[1])
    with torch.autocast("cuda", dtype=torch.bfloat16):
        hid = F.gelu(pc.ffn1[0](corr))
    print("ffn1[2] 1x1 54->36 ", variants(pc.ffn1[2], hid.float())[1])
    with torch.autocast("cuda", dtype=torch.bfloat16):
        f = pc.ffn1(corr)
        s = corr + f
        x1 = F.gelu(s)
    print("dtypes: ffn1", f.dtype, "x+ffn1", s.dtype, "gelu", x1.dtype)
    print("x1 == gelu_fp32(corr + float(ffn1))", float((x1 == F.gelu(corr + f.float())).float().mean()))
    print("dw k=1 groups=36   ", variants(pc.conv_list[0], x1, groups=36)[1])
    print("dw k=7 groups=36   ", variants(pc.conv_list[1], x1, groups=36)[1])
    print("pw 1x1 36->36      ", variants(pc.pw, x1)[1])
    with torch.autocast("cuda", dtype=torch.bfloat16):
        h2 = F.gelu(pc.ffn2[0](x1))
    print("ffn2[2] 1x1 54->256", variants(pc.ffn2[2], h2.float())[1])
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = F.gelu(pc(corr))
    print("block out dtype", out.dtype)
    # gelu of fp32 tensor: erf form in fp32?
    xs = torch.randn(100000, generator=g, device="cuda") * 3
    print("gelu fp32 == 0.5x(1+erf(x/sqrt2)) fp32:", float((F.gelu(xs) == 0.5 * xs * (1 + torch.erf(xs * 0.7071067811865476))).float().mean()))
