"""AdaptiveSoftmax.get_log_prob through the engine (tnt_decoder_log_prob: fp32 [rows, 50265], leading dimension 50265 -- no
16-byte-aligned rows) at the teacher-forced (2048 rows) and the decode-step (256 rows) size, plus pointer generation."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402
from tools.ab_dynamic import timeit  # noqa: E402

sd = random_state_dict(seed=0)
dec = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
for rows in (2048, 256):
    X = torch.randn(rows // 32, 32, 1024, device='cuda').bfloat16()
    ms = timeit(lambda: dec.get_normalized_probs((X, None)), 20, 3)
    print('get_log_prob rows=%d: %.3f ms (%.0f GB/s of fp32 output)' % (rows, ms, rows * 50265 * 4 / ms / 1e6))
