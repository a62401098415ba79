import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
print(sys.argv[1], [(r[ki][:14], r[vi]) for r in rows[1:]])
