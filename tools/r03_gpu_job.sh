set +e
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r03_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r03_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r03_bench.json 2> gpurun_out/r03_bench.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r03_ref_stages.csv python tools/ref_stage_table.py run c2 > gpurun_out/r03_ref_stages.log 2>&1; echo "ref stages rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r03_launches.csv python bench.py --steps 2 --warmup 1 --no-extras --no-cpu-baseline > gpurun_out/r03_ncu_bench.log 2>&1; echo "launch list rc=$?"
for tool in memcheck synccheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 python tools/sanitize.py > gpurun_out/r03_san_$tool.log 2>&1; echo "$tool rc=$?"; tail -4 gpurun_out/r03_san_$tool.log
done
timeout 900 compute-sanitizer --tool initcheck --error-exitcode 9 python tools/sanitize.py > gpurun_out/r03_san_initcheck.log 2>&1; echo "initcheck rc=$?"; tail -4 gpurun_out/r03_san_initcheck.log
