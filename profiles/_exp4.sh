TMD_UMMA_TRACE=1 TMD_NO_GRAPH=1 timeout 300 python profiles/_one_eval.py > gpurun_out/e4_trace.txt 2>&1; grep -c "umma trace" gpurun_out/e4_trace.txt
