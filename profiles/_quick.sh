# quick GPU check: conv + denoiser parity tests, then a short bench with the per-layer table
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/ev_q.csv > gpurun_out/b_q.json 2>gpurun_out/b_q.err || tail -5 gpurun_out/b_q.err
python -c "
import json;d=json.load(open('gpurun_out/b_q.json'));print('patches/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:(v['ms'],v['tflops']) for k,v in d['kernel_kinds'].items()})"
python profiles/layer_table.py gpurun_out/ev_q.csv 25
