import csv,sys
for f in sys.argv[1:]:
    rows=[r for r in csv.reader(open(f)) if len(r)>5]
    h=rows[0]; ik=h.index("Kernel Name"); iv=h.index("Metric Value")
    d={}
    for r in rows[1:]:
        if 'fxcu' in r[ik]: d.setdefault(r[ik][:70],[]).append(float(r[iv].replace(',',''))/1000)
    print(f)
    for k,v in d.items(): print("   ",k, [round(x,1) for x in v])
