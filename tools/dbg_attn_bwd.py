import os, sys, torch
dbg = torch.zeros(3 * 128, dtype=torch.int64, device="cuda")
os.environ["GXA_DBG_PTR"] = str(dbg.data_ptr())
sys.argv = ["x"]; sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests")); __file__ = os.path.join(os.getcwd(), "tests/bench_kernels.py")
exec(open("tests/bench_kernels.py").read())
torch.cuda.synchronize()
d = dbg.cpu().view(3, 16, 8)
t0 = int(d[d > 0].min())
for g in range(12):
    print(f'      CMP detail: top {d[1,g,5]-t0:7d} start {d[1,g,0]-t0:7d} sdpfull {d[1,g,1]-t0:7d} sdpfree-arrived {d[1,g,6]-t0:7d} math {d[1,g,2]-t0:7d} done {d[1,g,3]-t0:7d} loads-landed {d[1,g,4]-t0:7d}')
    print(f"g={g:2d} MMA: preready {d[0,g,0]-t0:7d} ready {d[0,g,1]-t0:7d} dqfree {d[0,g,2]-t0:7d} dkvfree {d[0,g,3]-t0:7d} issued {d[0,g,4]-t0:7d} | CMP: start {d[1,g,0]-t0:7d} sdpfull {d[1,g,1]-t0:7d} math {d[1,g,2]-t0:7d} done {d[1,g,3]-t0:7d} | DRN: start {d[2,g,0]-t0:7d} dqfull {d[2,g,1]-t0:7d}")
