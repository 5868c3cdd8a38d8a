"""Hand-off timeline of k_tc_fused_fwd (pair 0): python tools/fused_trace.py [T]  (C1 shape by default)"""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import inr_for_audio_b200 as P
from inr_for_audio_b200 import _lib as L
T = int(sys.argv[1]) if len(sys.argv) > 1 else 264600
torch.manual_seed(0)
m = P.Siren(1, 256, 3, 1, True)
gt = torch.from_numpy(P.utils.synth_waveform(T, seed=0)).cuda()
f = P.SirenFitter(m.spec, m.flat_parameters(), gt, precision="fp16", use_graph=False)
f.step(3)
L.check(L.lib().inr_dbg_fused_trace_enable(1))
f.step(1)
buf = (C.c_longlong * (96 * 4))()
L.check(L.lib().inr_dbg_fused_trace_read(buf, 96 * 4))
a = np.array(buf[:], dtype=np.int64).reshape(96, 4)
a = a[(a > 0).all(axis=1)]
t0 = a[0, 0]
print("job slot layer | mma_start mma_commit epi_start epi_end (clk from first) | mma_issue  commit->epi  epi_len  epi_end->next_mma_same_slot")
for j in range(min(len(a), 36)):
    b, l = j % 2, (j // 2) % 3 + 1
    nxt = a[j + 2, 0] - a[j, 3] if j + 2 < len(a) and l < 3 else -1
    print("%3d  %d  %d | %8d %8d %8d %8d | %6d %6d %6d %6d" % (j, b, l, *(a[j] - t0), a[j, 1] - a[j, 0], a[j, 2] - a[j, 1], a[j, 3] - a[j, 2], nxt))
