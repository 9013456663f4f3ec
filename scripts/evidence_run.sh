# Round evidence on one B200: bench lines, launch list, one `ncu --set full` capture.
# Usage (under gpurun): bash scripts/evidence_run.sh TAG
set -x
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
python bench.py > $O/${TAG}_bench_config5.json 2> $O/${TAG}_bench_config5.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_config5_reference_arm.json 2> $O/${TAG}_ref.err; echo "ref exit $?"
for w in config1 config2 config3 config4; do
  python bench.py --workload $w > $O/${TAG}_bench_$w.json 2> $O/${TAG}_bench_$w.err; echo "$w exit $?"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-validate --no-consumers --no-e2e-python --e2e-depth 1 > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu list exit $?"
ncu --set full --clock-control none --import-source on \
  -k regex:"k5_emit|k4_resolve|rs_onesweep|k1_count|k1_spine|p1_sites|k2_place|rs_global" -s 24 -c 24 \
  -o $O/prof_${TAG} -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-validate --no-consumers --no-e2e-python --e2e-depth 1 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la $O | tail -20
