#!/usr/bin/env python3
"""What the GPU box offers for the two third-party arithmetic cores of the reference (edlib, pyahocorasick; SURVEY.md
§8(c)) and for the host side: importable modules, cores, memory, GPUs.  Prints one JSON object."""
import importlib, json, os, platform, subprocess
out = {"python": platform.python_version(), "cpu_count": os.cpu_count()}
for m in ("edlib", "ahocorasick", "pysam", "matplotlib", "parasail", "Bio", "pyximport"):
    try:
        mod = importlib.import_module(m)
        out[m] = getattr(mod, "__version__", "importable")
    except Exception as e:
        out[m] = f"absent ({type(e).__name__})"
try:
    out["mem_total_gb"] = round(int(open("/proc/meminfo").readline().split()[1]) / 2**20, 1)
    out["cpu_model"] = [l.split(":")[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    out["numa_nodes"] = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
except Exception as e:
    out["hostinfo_error"] = str(e)
try:
    out["gpus"] = subprocess.run(["nvidia-smi", "--query-gpu=name,pcie.link.gen.max,pcie.link.width.max", "--format=csv,noheader"],
                                 capture_output=True, text=True).stdout.strip().splitlines()
except Exception as e:
    out["gpus"] = str(e)
for cmd in ("pip download edlib 2>&1 | tail -1",):
    pass
print(json.dumps(out))
