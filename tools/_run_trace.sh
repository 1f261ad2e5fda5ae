for st in 0 2 3; do for sh in 0 3; do timeout 120 python tools/tc_trace.py $sh $st 2>&1 | head -16; done; done > gpurun_out/trace_cta_s17.txt
cat gpurun_out/trace_cta_s17.txt
