# 2-GPU box: is the slow rank of the soup cast the DEVICE or the rays?  Same seed on both devices, then seeds swapped.
set -x; mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total,ecc.mode.current,retired_pages.pending,clocks.mem --format=csv
SOUP_SEED=0 timeout 600 python scripts/soup_ranks.py 2 > gpurun_out/soup_ranks_same_seed.log 2>&1; cut -c1-500 gpurun_out/soup_ranks_same_seed.log
timeout 600 python scripts/soup_ranks.py 2 1e7 4194304 > gpurun_out/soup_ranks_4m.log 2>&1; cut -c1-500 gpurun_out/soup_ranks_4m.log
