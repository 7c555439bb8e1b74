#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
./tools/diag/l2_sector_probe > gpurun_out/probe_plain.log 2>&1
cat gpurun_out/probe_plain.log
ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none --csv --log-file gpurun_out/probe_ncu.csv ./tools/diag/l2_sector_probe > gpurun_out/probe_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/probe_ncu.csv')) if len(r)>10]
h=rows[0]
ik=h.index('Kernel Name'); im=h.index('Metric Name'); iv=h.index('Metric Value'); ii=h.index('ID')
for r in rows[1:]:
    print(r[ii], r[ik][:20], r[im], r[iv])
PY
