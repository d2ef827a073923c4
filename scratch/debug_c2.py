import sys; sys.path.insert(0, '.')
import numpy as np
import llm_slo_ebpf_toolkit_b200 as pkg
from llm_slo_ebpf_toolkit_b200 import synth, _abi as A
from oracle import cpu as O
w = synth.generate_host(synth.c2())
ctx = pkg.Context(0)
g = ctx.process_batch(w['spans'], w['events'], w['idents'])
o = O.process_batch(w['spans'], w['events'], w['idents'], indexed=True, nthreads=16)
print('gpu cand', int(g['n_cand'].sum()), 'oracle', int(o['n_cand'].sum()), g['debug'].as_tuple(), o['debug'].as_tuple())
bad = np.nonzero(g['n_cand'] != o['n_cand'])[0]
print('spans with different n_cand:', bad[:20], len(bad))
badc = np.nonzero((g['cand_event'] != o['cand_event']).any(1))[0]
print('spans with different cand_event:', badc[:20], len(badc))
for i in bad[:5]:
    s = w['spans'][i]
    print('span', i, s, 'grp', w['idents']['grp'][s['ident']], 'gpu n', g['n_cand'][i], 'oracle n', o['n_cand'][i], g['cand_event'][i], o['cand_event'][i])
    ev = w['events']
    grp = w['idents']['grp'][ev['ident']]
    m = (grp == w['idents']['grp'][s['ident']]) & (np.abs(ev['ts_ns'] - s['ts_ns']) <= 2_000_000_000)
    idx = np.nonzero(m)[0]
    print(' events in window:', len(idx), 'min dt', (ev['ts_ns'][idx]-s['ts_ns']).min(), 'max dt', (ev['ts_ns'][idx]-s['ts_ns']).max())
    tmin = ev['ts_ns'].min()
    print(' span q', (s['ts_ns']-tmin)>>24, ' span ts-tmin', s['ts_ns']-tmin)
