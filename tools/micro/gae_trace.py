"""Per-CTA phase stamps of gae_tma_kernel at C3 (developer build, tools/micro/build_trace.sh)."""
import ctypes as C, os, sys
import numpy as np, torch
here = os.path.dirname(os.path.abspath(__file__))
libname = sys.argv[2] if len(sys.argv) > 2 else "libb200xp_trace.so"
lib = C.CDLL(os.path.join(here, "..", "..", "agilerl_b200", "_lib", "libb200xp.so") if libname == "product"
             else os.path.join(here, "_trace", libname))
TRACE = hasattr(lib, "xp_gae_trace_read")
vp, i64, f64 = C.c_void_p, C.c_int64, C.c_double
lib.xp_gae.restype = C.c_int
lib.xp_gae.argtypes = [vp] * 7 + [i64] * 4 + [f64, f64, vp]
if TRACE: lib.xp_gae_trace_read.argtypes = [vp, C.c_int]
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(4)
T, N, PP = 128, 1024, int(sys.argv[1]) if len(sys.argv) > 1 else 16
sets = []
SLAB = os.environ.get("XP_TRACE_SLAB", "0") == "1"   # values right behind rewards, returns behind advantages
for _ in range(8):
    if SLAB:
        rv = torch.randn(2, PP, T, N, device=dev, generator=g); r, v = rv[0], rv[1]
    else:
        r = torch.randn(PP, T, N, device=dev, generator=g); v = torch.randn(PP, T, N, device=dev, generator=g)
    d = torch.rand(PP, T, N, device=dev, generator=g) < 0.02
    lv = torch.randn(PP, N, device=dev, generator=g).double(); ld = torch.zeros(PP, N, device=dev, dtype=torch.float64)
    ag = torch.empty(2, PP, T, N, device=dev)
    sets.append((r, v, d, lv, ld, ag[0], ag[1]) if SLAB else (r, v, d, lv, ld, torch.empty_like(r), torch.empty_like(r)))
st = torch.cuda.current_stream().cuda_stream
def run(s):
    r, v, d, lv, ld, a, gg = s
    rc = lib.xp_gae(r.data_ptr(), v.data_ptr(), d.data_ptr(), lv.data_ptr(), ld.data_ptr(), a.data_ptr(), gg.data_ptr(),
                    T, N, PP, T * N, 0.99, 0.95, st)
    assert rc == 0, rc
for s in sets: run(s)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
for k in range(40): run(sets[k % 8])
ev[1].record(); torch.cuda.synchronize()
side = torch.cuda.Stream()
gr = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    st_save, st = st, side.cuda_stream
    for s_ in sets: run(s_)
    side.synchronize()
    with torch.cuda.graph(gr, stream=side):
        for k in range(40): run(sets[k % 8])
    st = st_save
gr.replay(); torch.cuda.synchronize()
best = 1e9
for rep in range(7):
    ev[0].record(); gr.replay(); ev[1].record(); torch.cuda.synchronize()
    best = min(best, ev[0].elapsed_time(ev[1]) * 25)
print(f"{libname} XP_GAE_KERNEL={os.environ.get('XP_GAE_KERNEL', '')} slab={int(SLAB)} pair={os.environ.get('XP_GAE_PAIR', '')} P={PP}: us per launch (graph of 40, 8 rotating sets): best of 7 {best:.2f}")
if not TRACE: sys.exit(0)
nct = PP * (N // 32)
for trial in range(1):
    run(sets[trial]); torch.cuda.synchronize()   # one isolated launch
    buf = np.zeros(nct * 16, dtype=np.uint64)
    assert lib.xp_gae_trace_read(buf.ctypes.data, buf.size) == 0
    t = buf.reshape(nct, 16).astype(np.int64)
    smid, g0, g1 = t[:, 0], t[:, 1], t[:, 11]
    c = t[:, 2:11] - t[:, 2:3]         # cycles since the CTA's start
    names = ["start", "issued", "box0 landed", "box0 computed", "box0 stored", "box1 landed", "box1 computed", "box1 stored", "end"]
    print(f"--- isolated launch {trial}: first CTA start -> last CTA end {(g1.max() - g0.min()) / 1e3:.2f} us; "
          f"CTA starts spread {(g0.max() - g0.min()) / 1e3:.2f} us; CTA life median {np.median(g1 - g0) / 1e3:.2f} max {(g1 - g0).max() / 1e3:.2f} us")
    for k, nm in enumerate(names):
        col = c[:, k]
        print(f"  {nm:14s} cycles since CTA start: min {col.min():6d} median {int(np.median(col)):6d} p90 {int(np.percentile(col, 90)):6d} max {col.max():6d}")
    per_sm = np.bincount(smid, minlength=148)
    print("  CTAs per SM: min", per_sm.min(), "max", per_sm.max(), "SMs used", (per_sm > 0).sum())
    # the late starters
    late = np.argsort(g0)[-8:]
    print("  latest CTA starts (us after first):", np.round((g0[late] - g0.min()) / 1e3, 2).tolist(), "ends", np.round((g1[late] - g0.min()) / 1e3, 2).tolist())
    # compute cycles per box by CTAs-per-SM
    comp0 = c[:, 3] - c[:, 2]; comp1 = c[:, 6] - c[:, 5]
    for n in sorted(set(per_sm[smid])):
        m = per_sm[smid] == n
        print(f"  SMs with {n} CTAs: box0 compute median {int(np.median(comp0[m]))} max {comp0[m].max()} cycles; box1 compute median {int(np.median(comp1[m]))} max {comp1[m].max()}; box0 landed median {int(np.median(c[m, 2]))}")
