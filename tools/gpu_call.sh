mkdir -p gpurun_out
for kb in 2048 3072 4096 6144 8192; do echo "chunk_kb $kb"; RTB_CHUNK_KB=$kb timeout -k 5 100 python tools/e2e_breakdown.py 2>&1 | head -1; done > gpurun_out/e2e_chunks.txt; cat gpurun_out/e2e_chunks.txt
