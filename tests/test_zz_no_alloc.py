"""tcb_yolo_nms_static: the allocation-free YOLOv5 postprocess entry point.

It exists for callers that are not allowed to allocate GPU memory — a plugin inside a process whose
host has already given all of HBM to its own pools, or a per-frame callback that cannot pay for the
device-wide synchronisation of cudaMalloc — so its scratch is a fixed device array inside the library
instead of a context workspace. This file is named to be collected last: a run that stops at the first
failure has then already run every other test. It passes alone in a fresh process (the first call on a
device performs the one-time set-up) and after the other tests.

Checked here, through the C ABI: two calls in a row give exactly what tcb_yolo_nms gives (bit-exact,
the standard of tests/test_gpu_yolo.py) — each later call is smaller than the first, so it runs over
the earlier calls' left-overs in the shared array (one of them has more than 4096 candidates in a frame,
which is what reaches the array's sort area); sizes beyond the compiled-in maximum are refused.
"""
import numpy as np
import pytest
import torch

from triton_client_b200 import _lib, gpu, synth

pytestmark = pytest.mark.gpu


def _rows(out, cnt):
    out, cnt = out.cpu().numpy(), cnt.cpu().numpy()
    return [out[b, :cnt[b]] for b in range(out.shape[0])], cnt


def test_static_entry_matches_context_entry_on_two_calls_in_a_row(engine):
    rng = np.random.default_rng(8)
    dense = np.zeros((1, 9000, 7), np.float32)      # every row a candidate: more than 4096 per frame, so the ordering
    dense[0, :, 0:2] = rng.uniform(100, 400, (9000, 2))   # leaves shared memory and sorts in the static array's sort area
    dense[0, :, 2:4] = rng.uniform(20, 90, (9000, 2))
    dense[0, :, 4] = rng.uniform(0.5, 1, 9000)
    dense[0, :, 5:] = rng.uniform(0.6, 1, (9000, 2))
    calls = [
        (synth.yolo_output(4, 25200, 2, 4242), dict(conf_thres=0.3)),                                   # largest footprint in the array
        (dense, dict(conf_thres=0.25)),
        (synth.yolo_output(2, 3000, 5, 4243, objects=25, rows_per_object=20, tie_frac=0.5), dict(conf_thres=0.25, iou_thres=0.6, classes=[0, 2, 4], max_det=50)),
    ]
    for pred, kw in calls:  # back to back, no reset in between
        dev = torch.from_numpy(pred).cuda()
        got = gpu.yolo_nms_static(dev, **kw)
        want = engine.yolo_nms(dev, **kw)
        torch.cuda.synchronize()
        g_rows, g_cnt = _rows(got[0], got[1])
        w_rows, w_cnt = _rows(want[0], want[1])
        assert np.array_equal(g_cnt, w_cnt)
        assert sum(int(c) for c in g_cnt) > 0
        for a, b in zip(g_rows, w_rows):
            assert np.array_equal(a, b)
        assert torch.equal(got[2], want[2])  # rows with obj > conf_thres


def test_static_entry_rejects_sizes_beyond_its_scratch():
    l = _lib.lib()
    assert l.tcb_yolo_nms_static_scratch_bytes() < 16 << 20  # "modest": every GPU that loads the library holds it
    too_many = torch.zeros((8, 25200, 7), dtype=torch.float32, device="cuda")  # 201600 rows > 131072
    with pytest.raises(_lib.TcbError) as ei:
        gpu.yolo_nms_static(too_many, conf_thres=0.3)
    assert ei.value.code == -5 and "exceeds the static scratch" in str(ei.value)
    wide = torch.zeros((65, 8, 7), dtype=torch.float32, device="cuda")       # b > 64
    with pytest.raises(_lib.TcbError):
        gpu.yolo_nms_static(wide, conf_thres=0.3)
    ok = gpu.yolo_nms_static(torch.zeros((1, 8, 7), dtype=torch.float32, device="cuda"), conf_thres=0.3)  # still usable
    torch.cuda.synchronize()
    assert int(ok[1][0]) == 0
