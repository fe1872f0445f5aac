timeout 300 python -m pytest tests/test_gpu_ops.py -q -x -k modulated_conv3d 2>&1 | tail -2
timeout 300 python -m pytest tests/test_gpu_denoiser.py -q -x 2>&1 | tail -2
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/ev_q.csv > gpurun_out/b_q.json 2>gpurun_out/b_q.err || tail -5 gpurun_out/b_q.err
python -c "
import json;d=json.load(open('gpurun_out/b_q.json'));print(round(d['value'],1), {k:(v['ms'],v['tflops']) for k,v in d['kernel_kinds'].items()})"
