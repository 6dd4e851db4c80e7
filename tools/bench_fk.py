"""K5 timing study (GPU box): dm_fk_reward through the C ABI with preallocated outputs, several sizes, with and
without an L2 flush; prints every repetition so that outliers are visible."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deepmimic_diffusion_mujoco_b200 import _lib

H, T = 64, 35
L = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for N in [int(a) for a in sys.argv[1:]] or [1024, 8192, 16384, 65536]:
    x = torch.randn(N, H, T, device="cuda") * 0.3
    x[:, :, 2] += 0.9; x[:, :, 3] += 1.0
    ref = x[0].clone()
    reward = torch.empty(N, H, device="cuda"); sums = torch.zeros(8, device="cuda", dtype=torch.float64)
    body = torch.empty(N, H, 15, 3, device="cuda")
    for with_body in (False, True):
        for do_flush in (True, False):
            def call():
                _lib.check(L.dm_fk_reward(x.data_ptr(), ref.data_ptr(), N, H, ctypes.c_float(1 / 60),
                                          body.data_ptr() if with_body else None, reward.data_ptr(), sums.data_ptr(), _lib.stream_ptr()))
            call(); torch.cuda.synchronize()
            ts = []
            for _ in range(8):
                if do_flush: flush.zero_()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); call(); e1.record(); torch.cuda.synchronize()
                ts.append(round(e0.elapsed_time(e1) * 1e3, 1))
            b = N * H * (144 + (180 if with_body else 0))
            med = sorted(ts)[len(ts) // 2]
            print(json.dumps({"N": N, "body": with_body, "flush": do_flush, "us": ts, "GBps_median": round(b / med / 1e3, 1)}))
