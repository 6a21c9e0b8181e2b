import os, glob, subprocess
print("cpu_count", os.cpu_count(), "affinity", sorted(os.sched_getaffinity(0)))
for n in sorted(glob.glob("/sys/devices/system/node/node*")):
    try:
        print(n, open(n + "/cpulist").read().strip(), open(n + "/meminfo").read().split("\n")[0].strip())
    except OSError as e:
        print(n, e)
try:
    print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
except OSError as e:
    print(e)
try:
    import pynvml
    pynvml.nvmlInit()
    for i in range(pynvml.nvmlDeviceGetCount()):
        h = pynvml.nvmlDeviceGetHandleByIndex(i)
        n = (os.cpu_count() + 63) // 64
        print("gpu", i, pynvml.nvmlDeviceGetPciInfo(h).busId, [hex(x) for x in pynvml.nvmlDeviceGetCpuAffinity(h, max(n, 4))])
        try:
            print("   numa node of device:", open(f"/sys/bus/pci/devices/{pynvml.nvmlDeviceGetPciInfo(h).busId.lower()[4:]}/numa_node").read().strip())
        except OSError as e:
            print("   ", e)
except Exception as e:  # noqa: BLE001
    print("pynvml:", e)
print(subprocess.run("cat /proc/self/status | grep -i 'mems_allowed\\|cpus_allowed'; which numactl; cat /sys/fs/cgroup/cpuset.cpus.effective /sys/fs/cgroup/cpuset.mems.effective 2>/dev/null", shell=True, capture_output=True, text=True).stdout)
