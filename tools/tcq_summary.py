"""Summarise the tcq_gemm_kernel durations of a launch list made with tools/tcq_shapes.py (ncu --metrics gpu__time_duration.sum --csv)."""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value')
Ts = [int(a) for a in sys.argv[2:]] or [8, 32, 128]
names = ["qkv", "o", "gate_up", "down_q4", "down_q6", "lm_head/4"]
nbytes = [6144*4096*144/256, 4096*4096*144/256, 28672*4096*144/256, 4096*14336*144/256, 4096*14336*210/256, 32064*4096*210/256]
i = 0
for r in rows[1:]:
    if 'tcq_gemm_kernel' in r[ki]:
        ns = float(r[vi].replace(',', ''))
        print(f"T={Ts[i // 6]:3d} {names[i % 6]:10s} {ns / 1e3:8.1f} us  {nbytes[i % 6] / ns:7.1f} GB/s")
        i += 1
