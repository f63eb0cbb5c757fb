B="python bench.py --no-cpu --no-hamming --no-euroc --no-frontend --no-highres --batch 512 --distinct 64 --steps 2 --warmup 3 --sustain 0"
$B > gpurun_out/r2f_plain.json 2> gpurun_out/r2f_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:pyr_level -s 48 -c 8 -f -o gpurun_out/r2f_tma $B > gpurun_out/r2f_ncu1.log 2>&1
export ORBX_NO_TMA=1
$B > gpurun_out/r2f_plain2.json 2> gpurun_out/r2f_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:pyr_level -s 48 -c 8 -f -o gpurun_out/r2f_notma $B > gpurun_out/r2f_ncu2.log 2>&1
ls -la gpurun_out | grep r2f
