#!/bin/bash
# torchrun entry: rank 0 runs its script under ncu (launch list only), the other ranks run it plainly.
#   python -m torch.distributed.run ... --no-python tools/rank0_ncu.sh OUT.csv script.py args...
out=$1; shift
if [ "$LOCAL_RANK" = "0" ]; then
  exec ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file "$out" python "$@"
else
  exec python "$@"
fi
