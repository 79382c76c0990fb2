import sys, os, hashlib
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from trase_b200 import synth
sc = synth.make_scene(300000, 1352, 1014, 2)
h = {k: hashlib.md5(v.numpy().tobytes()).hexdigest()[:10] for k, v in sc.items() if torch.is_tensor(v)}
print(torch.get_num_threads(), h)
