"""Golden token ids from the UNMODIFIED reference tokenizers, executed here (the build container):

    python tests/golden/make_golden_tokenizer.py     # -> tests/golden/tokenizer_golden.json

  stablediffusionv2/ldm/models/clip_zh/simple_tokenizer.py   SimpleTokenizer + tokenize() on the shipped CLIP BPE vocab
  wukong-huahua/ldm/models/clip_zh/simple_tokenizer.py       WordpieceTokenizer on the shipped vocab_zh.txt, wrapped as
                                                             FrozenCLIPEmbedder_ZH.tokenize does (encoders/modules.py:31-50)
and both classes again on two SMALL vocabularies written by this script (tests/golden/mini_bpe.txt.gz,
tests/golden/mini_vocab_zh.txt), so that the CPU tests have a case that needs no file from the reference checkout.

`mindspore` is the torch-backed stand-in of ms_shim.py; `ftfy` (not installed) is replaced by the identity, so the
prompts are chosen to be text ftfy leaves alone (clean UTF-8, no curly quotes / full-width forms / entities).
"""
import gzip
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ms_shim  # noqa: E402

REF = "/root/reference/vision"

EN = ["a photograph of an astronaut riding a horse", "A painting of a squirrel eating a burger, highly detailed!",
      "an oil-painting of 3 cats' toys (don't touch) in 1920s style", "", "   multiple   spaces\tand\nnewlines  ",
      "naïve café déjà vu", "城市夜景 cyberpunk city at night", "hello<|endoftext|>world", "x" * 5 + " " + "word " * 90]
ZH = ["城市夜景 赛博朋克 格雷格·鲁特科夫斯基", "来自深渊 风景 绘画 写实风格", "一只戴着墨镜的猫, oil painting, 4K!", "",
      "Café Déjà-vu 2023年", "莫奈 撑阳伞的女人 月亮 梦幻", "山水画" * 40, "hello world unaffable"]


def load(path, name, pkg_stub=None):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    if pkg_stub:
        mod.__package__ = pkg_stub
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def write_mini_vocabs():
    """A few hundred merges taken as the most frequent adjacent pairs of a fixed word list (a toy BPE training), and a
    WordPiece vocabulary of characters + a few whole words / suffixes."""
    words = ("the a an of and in on at to for with painting photo photograph astronaut horse riding squirrel burger "
             "detailed highly cat cats dog city night light dark oil style hello world touch toys multiple spaces").split()
    corpus = [list(w[:-1]) + [w[-1] + "</w>"] for w in words for _ in range(3)]
    merges = []
    for _ in range(120):
        counts = {}
        for w in corpus:
            for p in zip(w, w[1:]):
                counts[p] = counts.get(p, 0) + 1
        if not counts:
            break
        best = max(sorted(counts), key=lambda p: counts[p])
        merges.append(best)
        new = []
        for w in corpus:
            out, i = [], 0
            while i < len(w):
                if i + 1 < len(w) and (w[i], w[i + 1]) == best:
                    out.append(w[i] + w[i + 1]); i += 2
                else:
                    out.append(w[i]); i += 1
            new.append(out)
        corpus = new
    with gzip.open(os.path.join(HERE, "mini_bpe.txt.gz"), "wb") as f:
        f.write(("#version: mini\n" + "\n".join(" ".join(m) for m in merges) + "\n").encode("utf-8"))
    zh = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + list("城市夜景赛博朋克风绘画写实格一只戴着墨镜的猫年月亮梦幻山水") + \
         list("abcdefghijklmnopqrstuvwxyz0123456789") + ["##" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"] + \
         ["hello", "world", "oil", "paint", "##ing", "##able", "un", "##aff", "cafe", "deja", "vu", ",", "!", "-", "·", "4k"]
    with open(os.path.join(HERE, "mini_vocab_zh.txt"), "w", encoding="utf-8") as f:
        f.write("\n".join(zh) + "\n")


def main():
    ms_shim.install()
    sys.modules["ftfy"] = types.SimpleNamespace(fix_text=lambda t: t)
    write_mini_vocabs()
    out = {"en_prompts": EN, "zh_prompts": ZH}

    sd2 = load(os.path.join(REF, "stablediffusionv2/ldm/models/clip_zh/simple_tokenizer.py"), "ref_sd2_tok")
    tk = sd2.SimpleTokenizer()
    out["sd2_encode"] = [tk.encode(p) for p in EN]
    out["sd2_tokenize"] = [np.asarray(sd2.tokenize([p], truncate=True).asnumpy() if hasattr(sd2.tokenize([p], truncate=True), "asnumpy")
                                      else sd2.tokenize([p], truncate=True)).astype(int).tolist()[0] for p in EN]
    out["sd2_decode"] = [tk.decode(ids) for ids in out["sd2_encode"]]
    mini = sd2.SimpleTokenizer(os.path.join(HERE, "mini_bpe.txt.gz"))
    out["mini_bpe_encode"] = [mini.encode(p) for p in EN[:5]]
    out["mini_bpe_vocab_size"] = len(mini.encoder)

    # wukong: the module does `from .utils import ...`; give it a package to resolve against
    pkg = types.ModuleType("ref_wk_pkg")
    pkg.__path__ = [os.path.join(REF, "wukong-huahua/ldm/models/clip_zh")]
    sys.modules["ref_wk_pkg"] = pkg
    load(os.path.join(REF, "wukong-huahua/ldm/models/clip_zh/utils.py"), "ref_wk_pkg.utils")
    wk = load(os.path.join(REF, "wukong-huahua/ldm/models/clip_zh/simple_tokenizer.py"), "ref_wk_pkg.simple_tokenizer",
              pkg_stub="ref_wk_pkg")

    def zh_tokenize(tokenizer, text, ctx=77):   # FrozenCLIPEmbedder_ZH.tokenize, encoders/modules.py:31-50
        sot, eot = tokenizer.encoder["[CLS]"], tokenizer.encoder["[SEP]"]
        ids = [sot] + tokenizer.encode(text) + [eot]
        if len(ids) > ctx:
            ids = ids[:ctx - 1] + [eot]
        return ids + [0] * (ctx - len(ids))

    wp = wk.WordpieceTokenizer()
    out["wk_pieces"] = [wp.text_tokenize(p) for p in ZH]
    out["wk_tokenize"] = [zh_tokenize(wp, p) for p in ZH]
    wpm = wk.WordpieceTokenizer(os.path.join(HERE, "mini_vocab_zh.txt"))
    out["mini_wp_pieces"] = [wpm.text_tokenize(p) for p in ZH]
    out["mini_wp_tokenize"] = [zh_tokenize(wpm, p) for p in ZH]
    bpe_wk = wk.BpeTokenizer()
    out["wk_bpe_encode"] = [bpe_wk.encode(p) for p in EN[:3]]
    with open(os.path.join(HERE, "tokenizer_golden.json"), "w", encoding="utf-8") as f:
        json.dump(out, f, ensure_ascii=False, indent=0)
    print({k: len(v) if hasattr(v, "__len__") else v for k, v in out.items()})


if __name__ == "__main__":
    main()
