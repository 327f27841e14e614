import sys, numpy as np
sys.path.insert(0, '/root/repo')
from quest_b200 import capi, circuits
gpu = capi.load_product(prec=1); cpu = capi.load(capi.ORACLE_LIB_F32)
ge = gpu.createQuESTEnv(); ce = cpu.createQuESTEnv()
import random
rng = random.Random(1)
nq = 5
names = circuits.DENSE_1Q + circuits.DIAG_1Q + circuits.CTRL_DENSE + circuits.CTRL_DIAG + circuits.MULTI
for fusion in (0, 1):
    gpu.lib.qb200_setFusion(fusion)
    for name in names:
        worst = 0
        for trial in range(6):
            ops = [o for o in circuits.random_circuit(nq, 400, trial + 10, measure=False, families=(name,))][:2]
            if not ops: continue
            st = []
            for lib, env in ((gpu, ge), (cpu, ce)):
                q = lib.createQureg(nq, env); circuits.prepare(lib, q, trial)
                circuits.run_circuit(lib, q, ops); st.append(lib.get_state(q)); lib.destroyQureg(q, env)
            a, b = st[0].view(np.float32), st[1].view(np.float32)
            worst = max(worst, int((a != b).sum()))
        print(f"fusion={fusion} {name:28s} mismatching components: {worst}")
