# The round-end evidence run (one GPU): GPU test-suite, smoke, both bench arms, ncu launch list and one
# ncu --set full capture of the dominant kernel.  gpurun --timeout 900 -- bash tools/evidence.sh ; results in gpurun_out/f_*
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/f_pytest.log
tail -4 gpurun_out/f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/f_smoke.log; tail -2 gpurun_out/f_smoke.log
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/f_bench_ref.json 2>> gpurun_out/f_bench.err; echo "ref exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 1 --warmup 1 > gpurun_out/f_ncu1.log 2>&1; echo "ncu1 exit $?"
ncu --set full --clock-control none --import-source on -k regex:k1_kmer_kernel -c 3 -o gpurun_out/f_kmer python bench.py --steps 1 --warmup 1 > gpurun_out/f_ncu2.log 2>&1; echo "ncu2 exit $?"
ncu -i gpurun_out/f_kmer.ncu-rep --page raw --csv > gpurun_out/f_kmer_raw.csv 2>/dev/null
ncu -i gpurun_out/f_kmer.ncu-rep --page details > gpurun_out/f_kmer_details.txt 2>/dev/null
ls -la gpurun_out | head -30
