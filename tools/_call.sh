mkdir -p gpurun_out
for n in 8 4; do
rm -rf /tmp/D$n; s=$(date +%s.%N); IBLB_ROOT=/tmp/D$n IBLB_GPUS=$n IBLB_YDIM=32768 IBLB_ITERATIONS=201 ./app 256 256 128 1.0 1.0 5 1.0 100 0 0 > gpurun_out/r02_app_c4_gpus$n.log 2>&1; echo "app$n rc=$? wall $(echo "$(date +%s.%N) - $s" | bc) s"; tail -n 3 gpurun_out/r02_app_c4_gpus$n.log; tail -n 4 /tmp/D$n/Raw/256/256/SimLog.txt; tail -n 2 /tmp/D$n/Flux/*-flux.dat; cp /tmp/D$n/Raw/256/256/SimLog.txt gpurun_out/r02_app_c4_gpus${n}_SimLog.txt
done
cmp /tmp/D8/Flux/256_256_128_1_1x5-flux.dat /tmp/D4/Flux/256_256_128_1_1x5-flux.dat && echo "flux files identical (8 vs 4 slabs)"
