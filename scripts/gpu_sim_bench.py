"""Development aid / measurement of SURVEY 8(f2): batched LPV_Model.sim roll-out, achieved HBM bandwidth."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from colaborativempc_b200 import scenarios as sc
from colaborativempc_b200.planner import BatchedPlannerLPV
from colaborativempc_b200.track import TrackMap

cfg, track = sc.shipped_config(), TrackMap("Highway")
N = 20
eng = BatchedPlannerLPV(*bench.planner_args(cfg, track, N), nn=0, max_batch=1)
S, A = 1024, 4
x0, xprev, uprev = sc.platoon_scenarios(S, A, N, cfg["dt"], track, seed0=0)
rep = int(sys.argv[1]) if len(sys.argv) > 1 else 64           # 4096 * 64 = 262144 models, 0.88 GB of traffic: larger than L2
xp = torch.from_numpy(np.ascontiguousarray(xprev.reshape(S * A, N + 1, 9))).cuda().repeat(rep, 1, 1).contiguous()
u = torch.from_numpy(np.ascontiguousarray(uprev.reshape(S * A, N, 2))).cuda().repeat(rep, 1, 1).contiguous()
B = xp.shape[0]
out = torch.empty_like(xp); st = torch.empty((B,), dtype=torch.int32, device="cuda")
for _ in range(3): eng.sim_batch(xp, u, out=out, status=st)
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
for a, b in ev:
    a.record(); eng.sim_batch(xp, u, out=out, status=st); b.record()
torch.cuda.synchronize()
ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
bytes_per = (N + 1) * 72 * 2 + N * 16 + 4
peak = 6556.8
try: peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
except Exception: pass
gbs = bytes_per * B / (ms * 1e-3) / 1e9
print(json.dumps({"kernel": "lpv_sim_kernel", "models": B, "horizon": N, "ms": ms, "models_per_s": B / (ms * 1e-3),
                  "algorithmic_bytes_per_model": bytes_per, "achieved_GBps": gbs, "peak_GBps": peak, "frac": gbs / peak,
                  "status_ok": int((st == 0).sum().item())}))
