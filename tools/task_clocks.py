"""Per-task (32-ray group) durations, node visits and the walk / candidate / rest cycle split of the forward launches.
Needs the library built with EGR_NVCC_EXTRA=-DEGR_TASK_CLOCKS (the kernels then write %globaltimer stamps and
counters into the tail of the hit log).  Usage: [SCENE=c4] python tools/task_clocks.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import egr_b200
from egr_b200 import scenes
C4 = os.environ.get("SCENE", "c2").lower() == "c4"  # SCENE=c4: the 3M anisotropic set (800x800 here as well)
N, R = (3_000_000 if C4 else 1_000_000), 800
dev = torch.device("cuda", 0)
scene = scenes.catacaustics_scene(N, seed=0) if C4 else scenes.kitchen_scene(N, seed=0)
cams = scenes.kitchen_cameras(100, R, R)
pc = scenes.SyntheticModel(scene, device=dev)
rtr = egr_b200.GaussianRaytracer(pc, R, R, ppll_forward_size=1000, ppll_backward_size=120_000_000)
rt = rtr.cuda_module
tg = scenes.random_targets(R, R, seed=1)
def pct(name, a):
    a = a[a > 0].astype(np.float64)
    if a.size == 0:
        print(name, "empty"); return
    q = np.percentile(a, [50, 90, 99, 99.9, 100])
    print(f"{name}: n={a.size} mean={a.mean():.1f} p50={q[0]:.0f} p90={q[1]:.0f} p99={q[2]:.0f} p99.9={q[3]:.0f} max={q[4]:.0f} sum={a.sum():.3g}")
for vi in (17,):
    view = scenes.SyntheticCamera(cams[vi], tg, device=dev)
    st = rt.get_stats()
    res = {}
    for nb in (0, 1):
        rt.get_config().num_bounces.fill_(nb)
        rt.get_metadata().total_num_calls.zero_()
        with torch.enable_grad():
            egr_b200.render(view, rtr)
        torch.cuda.synchronize()
        res[nb] = {k: getattr(st, k).cpu().numpy().astype(np.int64).ravel() for k in
                   ("num_traversed_per_pixel", "num_hits_per_pixel", "num_composited_per_pixel")}
    print("== view", vi)
    for k in res[0]:
        pct("step0 " + k, res[0][k]); pct("step1 " + k, res[1][k] - res[0][k])
    ps = rt._pixel_state
    hl = rt._hitlog  # the stamps live in the last 512 chunks of the hit log (egr_trace.cu, EGR_TASK_CLOCKS)
    CH = 128 + 16 * 32 * 8  # HITLOG_CHUNK_BYTES
    off = (hl.numel() // CH - 512) * CH
    ps = hl
    dbg2 = ps[off + 65536 * 16: off + 65536 * 16 + 2 * 32768 * 16].view(torch.int32).cpu().numpy().astype(np.int64).reshape(2, 32768, 4) & 0xFFFFFFFF
    dbg = ps[off: off + 2 * 32768 * 16].view(torch.int32).cpu().numpy().astype(np.int64).reshape(2, 32768, 4) & 0xFFFFFFFF
    for s, ng in ((0, 20000), (1, None)):
        d = dbg[s]
        if ng is None:
            ng = int((res[1]["num_traversed_per_pixel"] - res[0]["num_traversed_per_pixel"] > 0).sum() + 31) // 32
        t0 = d[:ng, 0]; dt = d[:ng, 1]
        t0 = np.where(t0 < t0.max() - (1 << 31), t0 + (1 << 32), t0)
        nmax = d[:ng, 2] & 0xFFFF; rmax = d[:ng, 2] >> 16; nsum = d[:ng, 3]
        if s == 1:
            e = dbg2[s][:ng]
            ec = e[:, 1] & 0x1FFFF
            print(f"   cycles: node-walk {e[:,0].sum()/e[:,2].sum():.2f}  candidates {ec.sum()/e[:,2].sum():.2f}  rest (select/composite/bounce) {1-(e[:,0].sum()+ec.sum())/e[:,2].sum():.2f}")
            outer = e[:, 3] & 0xFFF; surv = e[:, 3] >> 12  # outer rounds; sum over range-halves of the max-lane survivor count
            cyc_walk, cyc_cand, cyc_sph = e[:, 0] * 16.0, (e[:, 1] & 0x1FFFF) * 16.0, (e[:, 1] >> 17) * 64.0
            print(f"   warp-level outer rounds per task: mean={outer.mean():.1f} max={outer.max()}; max-lane survivors per round: {surv.sum()/max(outer.sum(),1):.1f}")
            print(f"   cycles per outer round: walk {cyc_walk.sum()/max(outer.sum(),1):.0f}  candidates {cyc_cand.sum()/max(outer.sum(),1):.0f} (of which sphere tests {cyc_sph.sum()/max(outer.sum(),1):.0f});  cycles per max-lane exact test {(cyc_cand.sum()-cyc_sph.sum())/max(surv.sum(),1):.0f}")
            print(f"   lane node visits per task: max-lane mean={nmax.mean():.0f} p99={np.percentile(nmax,99):.0f} max={nmax.max()}; lane-sum mean={nsum.mean():.0f}; leaf rounds max-lane mean={rmax.mean():.1f} max={rmax.max()}")
            cc = np.corrcoef(nmax, dt)[0,1]; print(f"   corr(dur, nmax)={cc:.2f}; ns per max-lane node visit: median={np.median(dt/np.maximum(nmax,1)):.0f}")
            order = np.argsort(-dt)[:5]
            print("   longest:", [(int(dt[i]/1e3), int(nmax[i]), int(rmax[i]), int(nsum[i])) for i in order])
        base = t0.min(); end = (t0 + dt).max()
        print(f"step{s} tasks={ng} span={(end-base)/1e3:.1f}us  task dur us: mean={dt.mean()/1e3:.1f} p50={np.percentile(dt,50)/1e3:.1f} p90={np.percentile(dt,90)/1e3:.1f} p99={np.percentile(dt,99)/1e3:.1f} max={dt.max()/1e3:.1f}")
        # how long after the last task START does the step end; and the busy fraction
        print(f"   last start at {(t0.max()-base)/1e3:.1f}us ; sum(dur)/(span*2960 slots)={dt.sum()/((end-base)*2960):.2f}")
        order = np.argsort(-dt)[:5]
        print("   longest tasks (idx, start us, dur us):", [(int(i), round((t0[i]-base)/1e3,1), round(dt[i]/1e3,1)) for i in order])
