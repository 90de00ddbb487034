import json,sys
for f in sys.argv[1:]:
    d=json.loads([l for l in open(f) if l.startswith('{')][-1])
    print(f, round(d['value']), d['unit'], round(d['ms_per_step'],2),'ms', d['clocks']['sm_mhz'], d.get('model_frac_of_sustained_bf16_peak'))
    for k,v in d['kernels'].items(): print('   %-16s %8.4f ms x%-3d share %.3f %s frac %.3f' % (k, v['ms_per_launch'], v['launches_per_step'], v['share_of_step'], v['bound'], v['frac']))
