"""Tokenizers, CPU tier: token ids bit-equal to tests/golden/tokenizer_golden.json, which
tests/golden/make_golden_tokenizer.py produced by executing the reference's own tokenizer classes.  The cases on the
two small vocabularies committed next to the goldens always run; the cases on the reference's shipped vocabularies run
where those files can be found (this container; $MDX_VOCAB_DIR elsewhere)."""
import json
import os

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def G():
    with open(os.path.join(GOLD, "tokenizer_golden.json"), encoding="utf-8") as f:
        return json.load(f)


def _have(name):
    from minddiffusion_b200 import tokenizer
    try:
        tokenizer.find_vocab(name)
        return True
    except FileNotFoundError:
        return False


def test_mini_bpe_matches_executed_reference(G):
    from minddiffusion_b200.tokenizer import SimpleTokenizer, tokenize
    tk = SimpleTokenizer(os.path.join(GOLD, "mini_bpe.txt.gz"))
    assert len(tk.encoder) == G["mini_bpe_vocab_size"]
    for prompt, want in zip(G["en_prompts"][:5], G["mini_bpe_encode"]):
        assert tk.encode(prompt) == want, prompt
        assert tk.decode(want).replace(" ", "") == "".join(prompt.lower().split())   # every pre-token ends in a space
    # byte alphabet: 256 distinct printable symbols, printable ASCII maps to itself, byte 0 moves to U+0100
    assert len(set(tk.byte_encoder.values())) == 256 and tk.byte_encoder[ord("a")] == "a"
    assert tk.byte_encoder[0] == chr(0x100)
    t = tokenize(["a cat", ""], context_length=8, tokenizer=tk)
    sot, eot = tk.encoder["<|startoftext|>"], tk.encoder["<|endoftext|>"]
    assert t.dtype == np.int64 and t.shape == (2, 8) and t[0, 0] == sot and t[1].tolist() == [sot, eot, 0, 0, 0, 0, 0, 0]
    with pytest.raises(RuntimeError, match="too long"):
        tokenize(["the city at night " * 4], context_length=8, tokenizer=tk)
    cut = tokenize(["the city at night " * 4], context_length=8, truncate=True, tokenizer=tk)
    assert cut[0, -1] == eot and cut[0, 0] == sot


def test_mini_wordpiece_matches_executed_reference(G):
    from minddiffusion_b200.tokenizer import WordpieceTokenizer, tokenize_zh
    tk = WordpieceTokenizer(os.path.join(GOLD, "mini_vocab_zh.txt"))
    for prompt, pieces, ids in zip(G["zh_prompts"], G["mini_wp_pieces"], G["mini_wp_tokenize"]):
        assert tk.text_tokenize(prompt) == pieces, prompt
        assert tokenize_zh([prompt], tokenizer=tk)[0].tolist() == ids, prompt
    assert "[UNK]" in tk.text_tokenize(chr(0x3A9))                      # no piece matches: the whole word is unknown
    assert tk.text_tokenize("unaffable") == ["un", "##aff", "##able"]   # greedy longest match first
    long = tokenize_zh([G["zh_prompts"][6]], tokenizer=tk)[0]
    assert long[-1] == tk.encoder["[SEP]"] and long[0] == tk.encoder["[CLS]"] and (long != 0).all()


@pytest.mark.skipif(not _have("bpe_simple_vocab_16e6.txt.gz"), reason="CLIP BPE vocabulary file not available")
def test_clip_bpe_on_shipped_vocab_matches_executed_reference(G):
    from minddiffusion_b200.tokenizer import SimpleTokenizer, tokenize
    tk = SimpleTokenizer()
    assert len(tk.encoder) == 49408 and tk.encoder["<|startoftext|>"] == 49406 and tk.encoder["<|endoftext|>"] == 49407
    for prompt, enc, full, dec in zip(G["en_prompts"], G["sd2_encode"], G["sd2_tokenize"], G["sd2_decode"]):
        assert tk.encode(prompt) == enc, prompt
        assert tokenize([prompt], truncate=True, tokenizer=tk)[0].tolist() == full
        assert tk.decode(enc) == dec
    for prompt, enc in zip(G["en_prompts"][:3], G["wk_bpe_encode"]):   # wukong's BpeTokenizer is the same algorithm
        assert tk.encode(prompt) == enc


@pytest.mark.skipif(not _have("vocab_zh.txt"), reason="WordPiece vocabulary file not available")
def test_wordpiece_on_shipped_vocab_matches_executed_reference(G):
    from minddiffusion_b200.tokenizer import WordpieceTokenizer, tokenize_zh
    tk = WordpieceTokenizer()
    for prompt, pieces, ids in zip(G["zh_prompts"], G["wk_pieces"], G["wk_tokenize"]):
        assert tk.text_tokenize(prompt) == pieces, prompt
        assert tokenize_zh([prompt], tokenizer=tk)[0].tolist() == ids


def test_text_cleanup_without_ftfy():
    from minddiffusion_b200 import tokenizer as T
    curly = chr(0x201C) + "hello" + chr(0x201D) + " " + "".join(chr(0xFF21 + i) for i in range(3)) + " " + chr(0xFB01) + "ne&amp;dandy\r\nx"
    assert T._fix_text(curly) == '"hello" ABC fine&dandy\nx'
    assert T.whitespace_clean("  a \t b\n\nc ") == "a b c"
    assert T._fix_text("e" + chr(0x301)) == chr(0xE9)                      # NFC


def test_glide_sentencepiece_helpers(tmp_path):
    """Encoder_SP semantics (chinese_tokenizer.py:95-143) on a SentencePiece model trained here on a toy corpus: the end
    token is the vocabulary size, prompts are cut / padded to text_ctx with a mask, and the generation batch is
    [prompt] * n + [empty] * n (caption_to_tokens.py:21-27)."""
    spm = pytest.importorskip("sentencepiece")
    from minddiffusion_b200 import tokenizer as T
    corpus = tmp_path / "corpus.txt"
    corpus.write_text("\n".join(["a painting of a cat", "a photo of a dog in the city", "the city at night",
                                 "one more line about painting dogs", "night city photo"] * 20), encoding="utf-8")
    spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=str(tmp_path / "toy"), vocab_size=30, hard_vocab_limit=False,
                                   model_type="unigram", character_coverage=1.0, minloglevel=2)
    enc = T.SentencePieceEncoder(str(tmp_path / "toy.model"))
    assert enc.n_vocab == enc.num_tokens + 1 and enc.end_token == enc.num_tokens
    ids = enc.encode("a painting of a cat")
    assert ids == enc.sp.EncodeAsIds("a painting of a cat") and enc.decode(ids + [enc.end_token]) == "a painting of a cat"
    toks, mask = enc.padded_tokens_and_mask(ids, 32)
    assert len(toks) == 32 and toks[:len(ids)] == ids and set(toks[len(ids):]) == {enc.end_token}
    assert mask == [True] * len(ids) + [False] * (32 - len(ids))
    toks, mask = enc.padded_tokens_and_mask(list(range(40)), 16)
    assert toks == list(range(16)) and all(mask)
    t, m = T.convert_input_to_token_gen("the city at night", 3, 32, enc)
    assert t.shape == (6, 32) and t.dtype == np.int32 and m.dtype == np.int32
    assert (t[3:] == enc.end_token).all() and (m[3:] == 0).all() and (t[0] == t[2]).all() and m[0].sum() > 0
    t2, m2 = T.convert_input_to_token_super_res("the city at night", 2, 32, enc)
    assert t2.shape == (2, 32) and (t2[0] == t[0]).all() and (m2[0] == m[0]).all()
