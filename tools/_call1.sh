set -x
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
(python tools/cli_e2e.py cfg2; FBG_HOST_WRITER=1 python tools/cli_e2e.py cfg2; python tools/cli_e2e.py cfg3 0.2; FBG_HOST_WRITER=1 python tools/cli_e2e.py cfg3 0.2) > gpurun_out/c1_cli.log 2>&1
cat gpurun_out/c1_cli.log
