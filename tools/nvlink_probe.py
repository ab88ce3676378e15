#!/usr/bin/env python
"""Which NVLink byte counters does this box expose?  (bench.py's nvlink_tx_kib tries NVML field values, then nvidia-smi)"""
import subprocess
try:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    for name in ("NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX", "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX", "NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX"):
        fid = getattr(pynvml, name, None)
        print(name, fid)
        if fid is None:
            continue
        for arg in ([(fid, 0xFFFFFFFF)], [fid], [(fid, 0)]):
            try:
                v = pynvml.nvmlDeviceGetFieldValues(h, arg)[0]
                print("  ", arg, "ret", v.nvmlReturn, "type", v.valueType, "ull", v.value.ullVal, "ul", v.value.ulVal)
            except Exception as e:
                print("  ", arg, "EXC", repr(e))
except Exception as e:
    print("pynvml:", repr(e))
for cmd in (["nvidia-smi", "nvlink", "-gt", "d", "-i", "0"], ["nvidia-smi", "nvlink", "-s", "-i", "0"]):
    r = subprocess.run(cmd, capture_output=True, text=True)
    print(" ".join(cmd), "rc", r.returncode, (r.stdout + r.stderr)[:600])
