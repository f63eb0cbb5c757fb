B="python bench.py --no-cpu --no-highres --batch 64 --distinct 16 --steps 3 --warmup 3 --sustain 0 --no-strong"
$B > gpurun_out/r2d_plain.json 2> gpurun_out/r2d_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:bf2nn -s 4 -c 2 -f -o gpurun_out/r2d_bf2nn $B > gpurun_out/r2d_ncu1.log 2>&1
ncu --set full --clock-control none -k 'regex:window_kernel|grid_build_kernel|scan_counts|bow_dist|stereo_match|stereo_rows|stereo_outlier' -s 20 -c 24 -f -o gpurun_out/r2d_percall $B > gpurun_out/r2d_ncu2.log 2>&1
ncu --set full --clock-control none -k 'regex:init_match|grid_build_batch|vocab_descend|distinctive|merge_slices' -c 10 -f -o gpurun_out/r2d_batch $B > gpurun_out/r2d_ncu3.log 2>&1
H="python bench.py --no-cpu --no-hamming --no-euroc --no-frontend --batch 16 --distinct 4 --steps 3 --warmup 3 --sustain 0"
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2d_highres_launches.csv $H > gpurun_out/r2d_ncu4.log 2>&1
ls -la gpurun_out | tail -12
