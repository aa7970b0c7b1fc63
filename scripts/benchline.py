import sys,json
for l in sys.stdin:
    if not l.startswith("{"): continue
    j=json.loads(l); k=j["roofline"]["kernels"]
    print(round(j["value"]/1e6,1), round(j["ms_per_step"],3), {a:round(b["ms"]/b["launches"]*1000,1) for a,b in k.items()})
