cd /root/repo
L=tensorflow-distributed-clustering_b200/libkmc_b200.so
cp $L /tmp/keep.so; cp build_variants/var_noepi.so $L
for t in 1 16; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tc_assign_kernel -c 30 --csv --log-file gpurun_out/abl_$t.csv python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --param 3=$t > gpurun_out/abl_$t.log 2>&1
  python - gpurun_out/abl_$t.csv $t <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10 and r[0].isdigit()]
print("terms", sys.argv[2], [round(float(r[-1])/1e3,1) for r in rows if float(r[-1])>20000][:8])
PY
done
cp /tmp/keep.so $L
