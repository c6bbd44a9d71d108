# Round 2 (1 GPU): the allocation-free K3 entry point (qemb_gram_noalloc): its test alone in a fresh process, then the
# whole GPU suite (its file is collected last), smoke, and a short bench if minutes remain.
set -x
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -s > gpurun_out/r2c17_noalloc_alone.log 2>&1; echo "noalloc alone exit $?"; tail -n 5 gpurun_out/r2c17_noalloc_alone.log
timeout 540 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2c17_pytest.log 2>&1; echo "pytest exit $?"; tail -n 14 gpurun_out/r2c17_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2c17_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/r2c17_smoke.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-file > gpurun_out/r2c17_bench.json 2> gpurun_out/r2c17_bench.err; echo "bench exit $?"; tail -c 600 gpurun_out/r2c17_bench.json
