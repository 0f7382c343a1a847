import json, sys
l=[x for x in open(sys.argv[1]) if x.startswith('{')][-1]
d=json.loads(l)
k=d.pop('kernels')
print('value', round(d['value'],1), d['unit'], 'ms/step', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'], 'clocks', d['clocks'])
print('roofline', d['roofline'])
print('cpu', d['cpu_baseline'] and d['cpu_baseline']['value'], 'model_roofline', d['model_roofline'])
for n,v in k.items(): print(f"{n:22s} {v['ms_per_step']:9.2f} ms {v['share']*100:5.1f}% {v['launches_per_step']:5d} launches {v['GBps']:8.0f} GB/s {v['TFLOPs']:7.2f} TF")
