B=bench/gather4_probe.bin
mkdir -p gpurun_out
{
nvidia-smi -L
timeout 60 $B check 1
timeout 60 $B check 4
for st in 4 8 12; do timeout 60 $B bw 1 0 $st 1 0 0; done
timeout 60 $B bw 1 0 8 2 0 0
timeout 60 $B bw 1 0 12 4 0 0
timeout 60 $B bw 1 0 8 2 1 0
timeout 60 $B bw 1 0 8 2 0 1
timeout 60 $B bw 1 0 8 2 1 1
for st in 4 8 12; do timeout 60 $B bw 2 2 $st 1 0 0; done
timeout 60 $B bw 2 2 8 2 0 0
timeout 60 $B bw 2 2 12 2 0 0
timeout 60 $B bw 2 2 12 4 0 0
timeout 60 $B bw 2 2 12 2 1 0
timeout 60 $B bw 2 2 8 2 0 1
timeout 60 $B bw 2 2 12 2 0 1
timeout 60 $B bw 2 2 12 2 1 1
timeout 60 $B bw 2 1 12 2 1 1
} 2>&1 | tee gpurun_out/g4_probe.txt
