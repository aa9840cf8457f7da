import csv, sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]
from collections import OrderedDict
d=OrderedDict()
for r in rows[1:]:
    k=(r[0], r[h.index('Kernel Name')].split('(')[0][-34:])
    d.setdefault(k,{})[r[h.index('Metric Name')]]=r[h.index('Metric Value')]
f=lambda s: float(s.replace(',',''))
for k,v in d.items():
    print('%3s %-34s ms=%8.2f rd=%7.1fGB wr=%6.1fGB l2hit=%5.1f inst=%6.2fG'%(k[0], k[1], f(v['gpu__time_duration.sum'])/1e6, f(v['dram__bytes_read.sum'])/1e9, f(v['dram__bytes_write.sum'])/1e9, f(v.get('lts__t_sector_hit_rate.pct','nan')), f(v['smsp__inst_executed.sum'])/1e9))
