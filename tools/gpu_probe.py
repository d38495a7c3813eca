"""First GPU trip: FP64/FP32 pipe peaks for the roofline denominators."""
import ctypes as C
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nonmyopia_b200 import _lib

os.makedirs("gpurun_out", exist_ok=True)
res = {"gpu": torch.cuda.get_device_name(0), "sm_count": _lib.lib().nmb_device_sm_count()}
for dt, nm in ((0, "f64"), (1, "f32")):
    for which, wn in ((0, "fma_flops"), (1, "sincospi_per_s"), (2, "exp_per_s")):
        out = C.c_double(0)
        _lib.check(_lib.lib().nmb_pipe_peak(dt, which, C.byref(out)), "peak")
        res[f"{nm}_{wn}"] = out.value
print(json.dumps(res, indent=1))
json.dump(res, open("gpurun_out/pipe_peaks.json", "w"), indent=1)
