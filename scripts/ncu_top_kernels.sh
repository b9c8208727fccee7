# usage: bash scripts/ncu_top_kernels.sh <kernel regex> <out name>
set -e
K=${1:-k_enc}
OUT=${2:-prof}
python bench.py --containers 64 --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"$K" -s 2 -c 1 -o gpurun_out/$OUT python bench.py --containers 64 --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
