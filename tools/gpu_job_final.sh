mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --config 4 > gpurun_out/r02_config4_n1.json 2> gpurun_out/r02_config4_n1.err; echo "config4 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/r02_config4_n1.json')); print(d['value'], d['text_gbs'], d.get('parity'), d.get('cpu_baseline'))"
timeout 600 python bench.py --config 4 --order baseline > gpurun_out/r02_config4_n1_baseline_order.json 2>> gpurun_out/r02_config4_n1.err; echo "config4 baseline rc=$?"
python -c "import json; d=json.load(open('gpurun_out/r02_config4_n1_baseline_order.json')); print(d['value'], d['text_gbs'], d.get('parity'))"
timeout 600 python bench.py --config 3 > gpurun_out/r02_config3.json 2> gpurun_out/r02_config3.err; echo "config3 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/r02_config3.json')); print(d['value'], d['text_gbs'], d.get('parity'), d.get('roofline'))"
timeout 300 python bench.py --config 1 > gpurun_out/r02_config1.json 2>/dev/null; timeout 300 python bench.py --config 1 --impl reference > gpurun_out/r02_config1_ref.json 2>/dev/null; python -c "
import json
for f in ['r02_config1','r02_config1_ref']:
    d=json.load(open('gpurun_out/%s.json'%f)); print(f, d['value'])"
