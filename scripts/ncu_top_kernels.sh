set -e
python bench.py --containers 64 --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_enc1|k_dec1$' -c 2 -o gpurun_out/prof_r1_v2 python bench.py --containers 64 --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
