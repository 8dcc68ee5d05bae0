"""Parameter-shape lists of the BASELINE.json configs (synthetic workloads; no model code, no weights).

Counts are asserted against SURVEY.md section 8: ResNet-18 62 tensors / 11,689,512 elements, ResNet-50 161 /
25,557,032, BERT-base 206 / 110,106,428, GPT-2 medium 292 / 354,823,168.
"""
from __future__ import annotations

import numpy as np


def _resnet(block: str, layers):
    shapes = [(64, 3, 7, 7), (64,), (64,)]
    inplanes = 64
    exp = 4 if block == "bottleneck" else 1
    for i, (planes, n) in enumerate(zip((64, 128, 256, 512), layers)):
        stride = 1 if i == 0 else 2
        for b in range(n):
            if block == "bottleneck":
                shapes += [(planes, inplanes, 1, 1), (planes,), (planes,), (planes, planes, 3, 3), (planes,), (planes,),
                           (planes * 4, planes, 1, 1), (planes * 4,), (planes * 4,)]
            else:
                shapes += [(planes, inplanes, 3, 3), (planes,), (planes,), (planes, planes, 3, 3), (planes,), (planes,)]
            if b == 0 and (stride != 1 or inplanes != planes * exp):
                shapes += [(planes * exp, inplanes, 1, 1), (planes * exp,), (planes * exp,)]
            inplanes = planes * exp
    shapes += [(1000, 512 * exp), (1000,)]
    return shapes


def _bert_base(vocab=30522, h=768, layers=12, ffn=3072, max_pos=512, type_vocab=2):
    s = [(vocab, h), (max_pos, h), (type_vocab, h), (h,), (h,)]
    for _ in range(layers):
        s += [(h, h), (h,)] * 4 + [(h,), (h,)] + [(h, ffn), (ffn,), (ffn, h), (h,)] + [(h,), (h,)]
    s += [(h, h), (h,)]                       # pooler
    s += [(h, h), (h,), (h,), (h,), (vocab,)]  # MLM transform + LN + decoder bias (decoder weight tied)
    s += [(h, 2), (2,)]                       # NSP
    return s


def _gpt2_medium(vocab=50257, h=1024, layers=24, ctx=1024):
    s = [(vocab, h), (ctx, h)]
    for _ in range(layers):
        s += [(h,), (h,), (h, 3 * h), (3 * h,), (h, h), (h,), (h,), (h,), (h, 4 * h), (4 * h,), (4 * h, h), (h,)]
    s += [(h,), (h,)]
    return s


_EXPECT = {"resnet18": (62, 11_689_512), "resnet50": (161, 25_557_032), "bert_base": (206, 110_106_428),
           "gpt2_medium": (292, 354_823_168)}


def param_shapes(name: str):
    if name == "resnet18":
        s = _resnet("basic", (2, 2, 2, 2))
    elif name == "resnet50":
        s = _resnet("bottleneck", (3, 4, 6, 3))
    elif name == "bert_base":
        s = _bert_base()
    elif name == "gpt2_medium":
        s = _gpt2_medium()
    elif name.startswith("flat"):  # flat<log2 n | 2e9>x<tensors>[r], e.g. flat26x1, flat2e9x1, flat26x1000r (ragged)
        logn, k = name[4:].split("x")
        ragged = k.endswith("r")
        k = int(k.rstrip("r"))
        n = 2_000_000_000 if logn == "2e9" else 1 << int(logn)
        if not ragged:
            return [(n // k,)] * k
        # ragged: log-uniform sizes over four decades (odd lengths, a few tiny tensors), same total
        rng = np.random.default_rng(7)
        wts = np.exp(rng.uniform(0.0, np.log(1e4), k))
        sizes = np.maximum(1, (wts / wts.sum() * n).astype(np.int64)) | 1
        sizes[:4] = (1, 7, 1000, 4097)
        sizes[-1] += n - int(sizes.sum())
        assert sizes.min() > 0 and int(sizes.sum()) == n
        return [(int(x),) for x in sizes]
    else:
        raise KeyError(name)
    nt, ne = _EXPECT[name]
    assert len(s) == nt and sum(int(np.prod(x)) for x in s) == ne, (name, len(s), sum(int(np.prod(x)) for x in s))
    return s


def param_numels(name: str):
    return [int(np.prod(s)) for s in param_shapes(name)]
