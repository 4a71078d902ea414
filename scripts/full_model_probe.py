"""GPU aid: whole-model frames/s and phase split, reference vs installed (oracle/full_model.py).
  python scripts/full_model_probe.py [frames H W kernel_size iters] [--no-reference]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import full_model

args = [a for a in sys.argv[1:] if not a.startswith("--")]
frames, H, W, ks, iters = (int(x) for x in args[:5]) if len(args) >= 5 else (40, 720, 1280, 20, 20)
res = full_model.measure(frames, H, W, ks, iters, reference="--no-reference" not in sys.argv)
print(json.dumps(res, indent=1))
