B="python bench.py --no-cpu --no-hamming --no-euroc --no-frontend --no-highres --batch 512 --distinct 64 --steps 1 --warmup 3 --sustain 0"
$B > gpurun_out/r2z_plain.json 2> gpurun_out/r2z_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:"pyr_level|border|cell_nms|quadtree|describe" -s 72 -c 24 -f -o gpurun_out/r2z_all $B > gpurun_out/r2z_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 72 -c 400 --csv --log-file gpurun_out/r2z_launch_list.csv $B > gpurun_out/r2z_ncu2.log 2>&1
H="python bench.py --no-cpu --no-highres --no-euroc --no-frontend --no-strong --batch 64 --distinct 16 --steps 2 --warmup 3 --sustain 0"
ncu --set full --clock-control none -k regex:bf2nn -s 4 -c 1 -f -o gpurun_out/r2z_bf2nn $H > gpurun_out/r2z_ncu3.log 2>&1
E="python bench.py --no-cpu --no-highres --no-hamming --no-frontend --batch 64 --distinct 16 --steps 2 --warmup 3 --sustain 0"
ncu --set full --clock-control none -k regex:"init_match|grid_build_batch" -s 2 -c 2 -f -o gpurun_out/r2z_init $E > gpurun_out/r2z_ncu4.log 2>&1
ls -la gpurun_out | grep r2n
