import sys
sys.path.insert(0,'/root/repo')
from tests import corpus
from replisense_rfq_b200 import Engine, build
from replisense_rfq_b200.engine import PinnedArena
build.build()
eng=Engine(0)
docs=[corpus.text_body(s) for s in range(100000)]
nb=sum(map(len,docs))
arena=PinnedArena(eng, sum((len(d)+15)&~15 for d in docs)+4096)
off,ln,_=arena.pack(docs)
eng.stage(arena,off,ln)
for _ in range(3): eng.text_run(); eng.fetch()
eng.set_profile(True)
eng.text_run(); eng.fetch()
print(nb/1e6,'MB', eng.last_run_ms(), eng.profile())
