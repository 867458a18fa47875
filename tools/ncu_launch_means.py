import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); gi=h.index("Grid Size")
seen={}
for r in rows[1:]:
    k=r[ki][:44]
    seen.setdefault(k,[]).append(float(r[vi])/1000)
for k,v in seen.items():
    if any(x in k for x in ("tile_","seg_sum","chunk_base","render_","preprocess","scan_","onesweep","duplicate","histogram")):
        print(k.ljust(46), len(v), round(sum(v)/len(v),1), "us")
