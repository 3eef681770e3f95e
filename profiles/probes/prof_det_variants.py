"""A/B of libtcb200 builds on the detectron post of bench.py (16 RetinaNet images): CUDA-event time per call.
Usage: python profiles/probes/prof_det_variants.py [lib.so ...]   (each library in its own process, TCB_LIB_PATH)"""
import os
import subprocess
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if os.environ.get("_DET_CHILD"):
    sys.path.insert(0, ROOT)
    import torch
    import bench
    from triton_client_b200.gpu import Engine
    dev = torch.device("cuda", 0)
    eng = Engine(0, workspace_bytes=3 << 30)
    wl = bench.make_workload(os.environ.get("_DET_WL", "detectron"), eng, dev, 0, 1)
    names = wl.stage_names
    for i in range(5):
        wl.step(i)
    torch.cuda.synchronize()
    tot = [0.0] * len(names)
    reps = 40
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)] for _ in range(reps)]
    for i in range(reps):          # queued back to back: the GPU runs behind the host, so the events time kernels, not launches
        wl.step(i, evs[i])
    torch.cuda.synchronize()
    for ev in evs[reps // 4:]:
        for k in range(len(names)):
            tot[k] += ev[k].elapsed_time(ev[k + 1]) * reps / (reps - reps // 4)
    print(os.environ.get("TCB_LIB_PATH", "default"), {n: round(t / reps, 4) for n, t in zip(names, tot)}, flush=True)
else:
    for lib in sys.argv[1:] or [""]:
        env = dict(os.environ, _DET_CHILD="1")
        if lib:
            env["TCB_LIB_PATH"] = os.path.abspath(lib)
        subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)
