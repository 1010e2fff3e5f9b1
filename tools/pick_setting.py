"""Reads a tools/ring_sweep.py report and prints shell `export` lines for the staged (file) setting
with the best load+scan rate, so that a session script can run bench.py / run_configs.py with it.
  eval "$(python tools/pick_setting.py gpurun_out/r02_ring_n8b.txt)" """
import re
import sys

best = None
for line in open(sys.argv[1]):
    m = re.match(r"(\S+)\s+pipeline-source\s+\S+\s+load\+scan\s+([\d.]+) GB/s", line)
    if not m:
        continue
    f = m.group(1).split(":")
    f += ["memcpy", "split", "file"][len(f) - 3:] if len(f) < 6 else []
    if f[5] != "file":
        continue
    if best is None or float(m.group(2)) > best[0]:
        best = (float(m.group(2)), f, m.group(1))
if best is None:
    sys.exit(0)
rate, f, text = best
out = [f"# best staged setting: {text} at {rate} GB/s"]
if int(f[0]) > 0:
    out.append(f"export NVME_STROM_READERS={f[0]}")
if int(f[1]) > 0:
    out.append(f"export NVME_STROM_SLOT_KB={f[1]}")
out.append(f"export NVME_STROM_SLOTS={f[2]}")
out.append(f"export NVME_STROM_COPY={f[3]}")
if f[4] == "none":
    out.append("export STROM_BENCH_NO_AFFINITY=1")
for extra in f[6:]:
    out.append(f"export {extra}")
print("\n".join(out))
