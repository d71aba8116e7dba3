import time, sys, numpy as np
sys.path.insert(0, '.')
t0=time.perf_counter()
from gfalign_b200 import capi, synth
t1=time.perf_counter(); print('import', round(t1-t0,3))
w=synth.make_workload('cfg3', A=1000000); t=w.tangle
nodes,_,_=synth.node_table(t)
po,ps=synth.candidate_paths(t, 64, seed=3)
t2=time.perf_counter()
e=capi.Engine(0); t3=time.perf_counter(); print('create', round(t3-t2,3))
e.set_alignments(w.offsets,w.steps,t.n_nodes); t4=time.perf_counter(); print('set_alignments', round(t4-t3,3))
e.set_universe(nodes); t5=time.perf_counter(); print('set_universe', round(t5-t4,3))
e.score_paths(po,ps,filter=True); t6=time.perf_counter(); print('first score', round(t6-t5,3))
e.score_paths(po,ps,filter=True); t7=time.perf_counter(); print('second score', round(t7-t6,4))
e.score_paths(po[:2],ps,filter=True); t8=time.perf_counter(); print('P=1 score', round(t8-t7,4))
e.close()
