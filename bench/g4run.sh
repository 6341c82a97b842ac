B=bench/gather4_probe.bin
mkdir -p gpurun_out
{
for st in 4 6 8; do timeout 30 $B bw 1 0 $st 1 0 2; done
timeout 30 $B bw 1 0 6 1 1 2
timeout 30 $B bw 1 0 8 1 1 2
for st in 4 8; do timeout 30 $B bw 2 2 $st 1 0 2; done
timeout 30 $B bw 2 2 8 1 1 2
timeout 30 $B bw 2 1 8 1 1 2
} 2>&1 | tee gpurun_out/g4_fwd_probe.txt
