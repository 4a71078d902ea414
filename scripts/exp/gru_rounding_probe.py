"""GPU aid: which rounding flow does torch/cuDNN follow for SKSepConvGRU3D under bf16 autocast?  Rebuilds the unit
from fp32 convolutions (TF32 off) on bf16-rounded operands with candidate rounding variants and reports the share of
outputs bit-identical to the module run under autocast -- per conv first, then for the whole unit."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn.functional as F
from oracle import ref_import
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
ref_import.load_reference()
upd = importlib.import_module("models.core.ppmtereo_update")
torch.manual_seed(0)
gru = upd.SKSepConvGRU3D(hidden_dim=128, input_dim=384).cuda().eval()
g = torch.Generator(device="cuda").manual_seed(1)
T, H, W = 4, 24, 40
h = torch.tanh(torch.randn(1, 128, T, H, W, generator=g, device="cuda")).bfloat16()
x = torch.randn(1, 384, T, H, W, generator=g, device="cuda")
r16 = lambda t: t.to(torch.bfloat16).float()
hx = torch.cat([h, x], 1)

def conv_variants(conv, inp):
    w, b = r16(conv.weight), conv.bias
    acc = F.conv3d(r16(inp).double(), w.double(), None, padding=conv.padding).float()      # "exact" accumulation
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ref = conv(inp).float()
    out = {"bf16(acc + bf16 bias)": r16(acc + r16(b).view(1, -1, 1, 1, 1)),
           "bf16(acc + fp32 bias)": r16(acc + b.view(1, -1, 1, 1, 1)),
           "bf16(bf16(acc) + bf16 bias)": r16(r16(acc) + r16(b).view(1, -1, 1, 1, 1))}
    return ref, out

with torch.no_grad():
    for name in ("convz1.0", "convz2", "convq3"):
        conv = gru
        for part in name.split("."):
            conv = conv[int(part)] if part.isdigit() else getattr(conv, part)
        ref, outs = conv_variants(conv, hx)
        print(name, {k: round(float((v == ref).float().mean()), 5) for k, v in outs.items()})
    # elementwise chain: feed torch's own bf16 conv outputs, vary only the gate arithmetic
    with torch.autocast("cuda", dtype=torch.bfloat16):
        z = torch.sigmoid(gru.convz2(hx)); r = torch.sigmoid(gru.convr2(hx))
        rh = r * h
        q = torch.tanh(gru.convq2(torch.cat([rh, x], 1)))
        ref = (1 - z) * h + z * q
        print("dtypes: z", z.dtype, "rh", rh.dtype, "cat", torch.cat([rh, x], 1).dtype, "q", q.dtype, "1-z", (1 - z).dtype, "out", ref.dtype)
    zf, hf, qf = z.float(), h.float(), q.float()
    cands = {"bf16(bf16(bf16(1-z)*h) + bf16(z*q))": r16(r16(r16(1 - zf) * hf) + r16(zf * qf)),
             "bf16((1-z)*h + z*q) fp32 chain": r16((1 - zf) * hf + zf * qf)}
    print({k: round(float((v == ref.float()).float().mean()), 5) for k, v in cands.items()})
    s = torch.sigmoid(gru.convz2(hx).float())      # sigmoid of the fp32-conv? no: check sigmoid rounding on bf16 input
    with torch.autocast("cuda", dtype=torch.bfloat16):
        c = gru.convz2(hx)
    print("sigmoid: bf16(1/(1+exp(-x)))", float((r16(1 / (1 + torch.exp(-c.float()))) == torch.sigmoid(c).float()).float().mean()),
          " gelu erf form", float((r16(0.5 * c.float() * (1 + torch.erf(c.float() * 0.7071067811865476))) == F.gelu(c).float()).float().mean()),
          " tanh", float((r16(torch.tanh(c.float())) == torch.tanh(c).float()).float().mean()))
