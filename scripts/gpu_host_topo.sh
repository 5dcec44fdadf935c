#!/bin/bash
# Host topology facts that bound the end-to-end (host array) numbers: NUMA nodes, the cpuset / memory nodes this
# container may use, where each GPU hangs.  Output: gpurun_out/r2_host_topo.log
out=gpurun_out/r2_host_topo.log
{
echo "== cpus"; nproc; grep -E "Cpus_allowed_list|Mems_allowed_list" /proc/self/status
echo "== numa nodes"; ls -d /sys/devices/system/node/node* 2>/dev/null
for n in /sys/devices/system/node/node*; do echo "$n cpulist=$(cat $n/cpulist) $(grep MemTotal $n/meminfo)"; done
echo "== gpus (pci numa_node, link)"
for d in /sys/bus/pci/devices/*; do
  if [ "$(cat $d/vendor)" = "0x10de" ] && [[ "$(cat $d/class)" == 0x0302* ]]; then
    echo "$(basename $d) numa=$(cat $d/numa_node) speed=$(cat $d/current_link_speed 2>/dev/null) width=$(cat $d/current_link_width 2>/dev/null) parent=$(basename $(dirname $(readlink -f $d)))"
  fi
done
echo "== nvidia-smi topo"; nvidia-smi topo -m 2>&1 | head -30
echo "== visible"; nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.width.current --format=csv
echo "== engine view"
python - <<'PY'
import ctypes as C, re
from fftw3_b200 import build
L = C.CDLL(build.lib_path("fp32"))
L.fftw_malloc.restype = C.c_void_p
L.fftw_malloc.argtypes = [C.c_size_t]
p = L.fftw_malloc(256 << 20)
C.memset(p, 1, 256 << 20)
for ln in open("/proc/self/numa_maps"):
    a = int(ln.split()[0], 16)
    if a <= p < a + (1 << 34) and re.search(r"N\d+=", ln) and abs(a - p) < (1 << 21):
        print("fftw_malloc(256 MiB) at %x:" % p, ln.strip()[:200])
PY
} > $out 2>&1
cat $out
