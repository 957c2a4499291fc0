import sys,json
for l in sys.stdin:
    if l.startswith('{"metric'):
        d=json.loads(l)
        print(d.get('impl'),'it/s',d['value'],'ms',d['ms_per_step'],'fwd_fps',d.get('fwd_fps'),'e2e',d.get('e2e',{}).get('value'))
        print('  '+' | '.join(f"{s['stage']} {s['ms']:.3f} ({s['frac']:.2f})" for s in d.get('stages',[])))
    elif l.strip(): print(l.rstrip()[:300])
