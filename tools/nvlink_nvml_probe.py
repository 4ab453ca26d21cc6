"""What NVML reports for the NVLink throughput fields on this box (run on >= 2 GPUs)."""
import pynvml, torch, time
pynvml.nvmlInit()
n = pynvml.nvmlDeviceGetCount()
def read(i):
    h = pynvml.nvmlDeviceGetHandleByIndex(i)
    out = {}
    for name in ("NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX", "NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX", "NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX", "NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX"):
        fid = getattr(pynvml, name)
        for scope in (0xFFFFFFFF, 0):
            try:
                t0 = time.perf_counter()
                v = pynvml.nvmlDeviceGetFieldValues(h, [(fid, scope)])[0]
                out[(name[-7:], scope)] = (v.nvmlReturn, v.valueType, int(v.value.ullVal), round((time.perf_counter() - t0) * 1e3, 2))
            except Exception as e:
                out[(name[-7:], scope)] = repr(e)
    try:
        out["links"] = [pynvml.nvmlDeviceGetNvLinkState(h, l) for l in range(18)]
    except Exception as e:
        out["links"] = repr(e)
    return out
print("gpus", n)
for i in range(n): print(i, read(i))
a = torch.ones(256 << 20, dtype=torch.uint8, device="cuda:0")
b = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:1")
for _ in range(10): b.copy_(a)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
print("after 10 x 256 MiB cuda:0 -> cuda:1")
for i in range(n): print(i, read(i))
