import json,subprocess,sys,os
wl=sys.argv[1] if len(sys.argv)>1 else 'fit5k'
out=subprocess.run([sys.executable,'bench.py','--workload',wl,'--steps','40','--no-cpu-baseline','--no-extras','--no-parity'],capture_output=True,text=True).stdout
d=json.loads(out)
k=d['roofline'].get('ms_per_kernel') or {'lbs':d['roofline']['ms_per_launch']}
print(wl,'value %.3f M'%(d['value']/1e6), {a:{x:round(y,4) for x,y in b.items() if y>0.03 or wl!='fit100k'} for a,b in k.items()} if isinstance(list(k.values())[0],dict) else k)
