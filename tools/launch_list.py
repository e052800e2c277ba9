import csv,sys
rows=list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
h=rows[0]; ki=h.index('Kernel Name'); vi=h.index('Metric Value'); gi=h.index('Grid Size'); bi=h.index('Block Size')
pat=sys.argv[2] if len(sys.argv)>2 else 'mc_'
for r in rows[1:]:
    if pat in r[ki]: print(r[ki][:60], r[gi], r[bi], r[vi])
