import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from oracle import loader
from ol_dsp_b200 import DattorroBank, dattorro
orc = loader.oracle(); NAN = float("nan")
SETTERS = [dattorro.PREDELAY, dattorro.PREFILTER, dattorro.INPUT_DIFFUSION1, dattorro.INPUT_DIFFUSION2, dattorro.DECAY, dattorro.DECAY_DIFFUSION, dattorro.DAMPING]
cases = {"default": None, "predelay0": [0.0]+[NAN]*6, "predelay.5": [0.5]+[NAN]*6, "predelay1": [1.0]+[NAN]*6, "prefilter.1": [NAN,0.1]+[NAN]*5,
         "indiff1": [NAN,NAN,0.5]+[NAN]*4, "indiff2": [NAN]*3+[0.5]+[NAN]*3, "decay.9": [NAN]*4+[0.9,NAN,NAN], "decay.5": [NAN]*4+[0.5,NAN,NAN],
         "decaydiff": [NAN]*5+[0.5,NAN], "damping": [NAN]*6+[0.5], "predelay.01": [0.01]+[NAN]*6}
names = list(cases); I = len(names); T = int(sys.argv[1]) if len(sys.argv) > 1 else 9000
chunks = [int(c) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [T]
x = orc.noise(5, 0, I, 0, T, 0.25)
b = DattorroBank(I)
for i, n in enumerate(names):
    p = cases[n]
    if p:
        for w, v in zip(SETTERS, p):
            if not np.isnan(v): b.Set(w, float(v), first=i, count=1)
dx = torch.from_numpy(x).cuda(); dl, dr = torch.empty_like(dx), torch.empty_like(dx)
a = 0
for n in chunks:
    b.ProcessDevice(dx[a:].data_ptr(), dl[a:].data_ptr(), dr[a:].data_ptr(), n); a += n
torch.cuda.synchronize(); print("kernel", b.LastKernel())
l, r = dl.cpu().numpy(), dr.cpu().numpy()
for i, n in enumerate(names):
    wl, wr = orc.dattorro(x[:, i], cases[n])
    bl = np.flatnonzero(l[:, i].view(np.uint32) != wl.view(np.uint32)); br = np.flatnonzero(r[:, i].view(np.uint32) != wr.view(np.uint32))
    print(f"{n:12s} L first mismatch {bl[0] if bl.size else None} ({bl.size}) R {br[0] if br.size else None} ({br.size})", (l[bl[0], i], wl[bl[0]]) if bl.size else "")
