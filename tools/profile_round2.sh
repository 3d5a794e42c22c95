set -x
B="python bench.py --steps 3 --warmup 3 --no-extras --no-e2e --no-cpu-baseline"
$B > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_oz_rows|k_oz_gemm' -s 8 -c 2 -o gpurun_out/r02_oz_full -f $B > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r02b.json 2> gpurun_out/bench_r02b.err; echo bench rc=$?
