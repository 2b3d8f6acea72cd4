import time, pynvml, torch
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
x = torch.zeros(1 << 28, device="cuda")
for name, fn in [("clock", lambda: pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)), ("reasons", lambda: pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)),
                 ("power", lambda: pynvml.nvmlDeviceGetPowerUsage(h))]:
    for busy in (False, True):
        ts = []
        for _ in range(5):
            if busy:
                for _ in range(50): x.add_(1.0)
            t = time.perf_counter(); v = fn(); ts.append((time.perf_counter() - t) * 1e3)
            torch.cuda.synchronize()
        print(name, "busy" if busy else "idle", ["%.2f" % t for t in ts], v)
# effect on sync latency
def sync_lat():
    t = time.perf_counter()
    for _ in range(200):
        x[:1024].add_(1.0); torch.cuda.synchronize()
    return (time.perf_counter() - t) / 200 * 1e3
print("sync latency ms (no sampler)", sync_lat())
import threading
stop = False
def run():
    while not stop:
        pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM); pynvml.nvmlDeviceGetCurrentClocksEventReasons(h); time.sleep(0.02)
th = threading.Thread(target=run); th.start()
print("sync latency ms (sampler 20ms)", sync_lat()); stop = True; th.join()
