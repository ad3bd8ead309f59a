import csv,sys
rows=[r for r in csv.reader(l for l in open('/root/repo/gpurun_out/launches.csv') if l.startswith('"'))]
hdr=rows[0]; rows=rows[1:]
ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); gi=hdr.index("Grid Size")
half=len(rows)//2
tot=0
for r in rows[half:]:
    print(r[ki].split('(')[0][-40:], r[gi], float(r[vi])/1000, 'us'); tot+=float(r[vi])/1000
print('total', tot)
