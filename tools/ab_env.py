"""Times the bench steps (fwd, gen, train; 1 GPU) in THIS process -- run it twice with different environment switches on the
same box to A/B a build-time-constant option (e.g. TNT_GEMM_EPI_SERIAL=1)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402
from news_image_caption_b200.trainer import Trainer  # noqa: E402
from tools.ab_dynamic import timeit  # noqa: E402

sd = random_state_dict(seed=0)
dec = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
prev, tgt, ctx = bench.make_inputs('fwd', 1234)
p, t = prev.cuda(), tgt.cuda()
c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
p2, t2, c2 = bench.make_inputs('gen', 1234)
c2 = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in c2.items()}
seed = p2[:, :1].contiguous().cuda()
res = []
for _ in range(3):
    res.append((timeit(lambda: dec.nll(dec(p, c)[0], t)), timeit(lambda: dec.generate(seed, c2, gen_len=32, trim=False), 5, 2)))
dect = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
tr = Trainer(dect, t_total=100000)
dect.set_dropout(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, seed=1)
trn = [timeit(lambda: tr.step(p, t, c), 20, 5) for _ in range(2)]
print(os.environ.get('AB_TAG', ''), 'fwd %s  gen %s  train %s' % (['%.4f' % r[0] for r in res], ['%.3f' % r[1] for r in res], ['%.3f' % x for x in trn]))
