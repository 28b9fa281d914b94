"""Prompt -> token ids for the three text towers.  Host code, pure Python.

  SimpleTokenizer / tokenize          stablediffusionv2: CLIP byte-level BPE, <|startoftext|> ... <|endoftext|>, zero
                                      padded to 77 (ldm/models/clip_zh/simple_tokenizer.py:74-169)
  WordpieceTokenizer / tokenize_zh    wukong-huahua: BERT basic tokenisation (CJK characters split out, lower-cased,
                                      accents stripped, punctuation split) + greedy longest-match WordPiece, [CLS] ...
                                      [SEP], cut to 77 (clip_zh/simple_tokenizer.py:180-312, encoders/modules.py:31-50)
  SentencePieceEncoder + helpers      Taichu-GLIDE: a SentencePiece model, end token = vocab size, padded to text_ctx
                                      with a mask (tokenizer/chinese_tokenizer.py:95-143, caption_to_tokens.py:21-41)

The vocabulary FILES are data the reference ships next to its tokenizers (`bpe_simple_vocab_16e6.txt.gz`,
`vocab_zh.txt`) or with its checkpoints (GLIDE's `cog-pretrain.model` is not in the repository at all); they are
not part of this package.  Every class takes the path explicitly; `find_vocab(name)` looks in $MDX_VOCAB_DIR, in
`<repo>/vocab/` and in a checkout of the reference next to this repository.

Text clean-up: the reference calls `ftfy.fix_text` first.  ftfy is used when importable; otherwise `_fix_text`
applies ftfy's deterministic normalisations (HTML entities, NFC, curly quotes, full-width Latin, ligatures, control
characters, line breaks) and does NOT attempt its mojibake repair — for text that is already clean UTF-8 the two
agree.
"""
import gzip
import html
import os
import unicodedata

import numpy as np
import regex

SOT, EOT = "<|startoftext|>", "<|endoftext|>"
N_MERGES = 49152 - 256 - 2        # simple_tokenizer.py:79

_VOCAB_DIRS = [os.environ.get("MDX_VOCAB_DIR", ""),
               os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "vocab"),
               "/root/reference/vision/stablediffusionv2/ldm/models/clip_zh",
               "/root/reference/vision/wukong-huahua/ldm/models/clip_zh"]


def find_vocab(name):
    for d in _VOCAB_DIRS:
        if d and os.path.exists(os.path.join(d, name)):
            return os.path.join(d, name)
    raise FileNotFoundError("%s not found; copy it from the reference checkout (ldm/models/clip_zh/) into a directory "
                            "named by $MDX_VOCAB_DIR" % name)


# ----------------------------------------------------------------------------------------------------------
# text clean-up
# ----------------------------------------------------------------------------------------------------------
_QUOTES = {0x2018: "'", 0x2019: "'", 0x201A: "'", 0x201B: "'", 0x201C: '"', 0x201D: '"', 0x201E: '"', 0x201F: '"'}
_LIGATURES = {0xFB00: "ff", 0xFB01: "fi", 0xFB02: "fl", 0xFB03: "ffi", 0xFB04: "ffl", 0xFB05: "ſt", 0xFB06: "st",
              0x0132: "IJ", 0x0133: "ij", 0x01C7: "LJ", 0x01C8: "Lj", 0x01C9: "lj", 0x01CA: "NJ", 0x01CB: "Nj",
              0x01CC: "nj", 0x01F1: "DZ", 0x01F2: "Dz", 0x01F3: "dz"}
_WIDTH = {cp: chr(cp - 0xFEE0) for cp in range(0xFF01, 0xFF5F)}
_WIDTH[0x3000] = " "
_TABLE = {}
_TABLE.update(_QUOTES)
_TABLE.update(_LIGATURES)
_TABLE.update(_WIDTH)
_TABLE.update({0x2028: "\n", 0x2029: "\n", 0x0085: "\n"})
_KEEP_CTRL = {"\t", "\n", "\r"}


def _fix_text(text):
    text = html.unescape(text)
    text = text.replace("\r\n", "\n").replace("\r", "\n")
    text = text.translate(_TABLE)
    text = "".join(ch for ch in text if ch in _KEEP_CTRL or unicodedata.category(ch) not in ("Cc", "Cs"))
    return unicodedata.normalize("NFC", text)


def basic_clean(text):
    """simple_tokenizer.py:62-65."""
    try:
        import ftfy
        text = ftfy.fix_text(text)
    except ImportError:
        text = _fix_text(text)
    return html.unescape(html.unescape(text)).strip()


def whitespace_clean(text):
    return regex.sub(r"\s+", " ", text).strip()


# ----------------------------------------------------------------------------------------------------------
# CLIP byte-level BPE
# ----------------------------------------------------------------------------------------------------------
def byte_alphabet():
    """The 256 byte values as printable unicode symbols: printable Latin-1 bytes stand for themselves, the other 68
    are moved to U+0100.. in byte order (simple_tokenizer.py:30-48)."""
    printable = list(range(0x21, 0x7F)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    table, spare = {}, 0
    for b in printable:
        table[b] = chr(b)
    for b in range(256):
        if b not in table:
            table[b] = chr(256 + spare)
            spare += 1
    return table, printable + [b for b in range(256) if b not in set(printable)]


class SimpleTokenizer(object):
    """`encode(text) -> [ids]`, `decode(ids) -> text`, `.encoder` / `.decoder` as the reference class."""

    PATTERN = regex.compile(r"<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[\p{L}]+|[\p{N}]|[^\s\p{L}\p{N}]+",
                            regex.IGNORECASE)

    def __init__(self, bpe_path=None):
        bpe_path = bpe_path or find_vocab("bpe_simple_vocab_16e6.txt.gz")
        self.byte_encoder, order = byte_alphabet()
        self.byte_decoder = {s: b for b, s in self.byte_encoder.items()}
        with gzip.open(bpe_path) as f:
            lines = f.read().decode("utf-8").split("\n")
        merges = [tuple(ln.split()) for ln in lines[1:N_MERGES + 1]]          # line 0 is the version header
        symbols = [self.byte_encoder[b] for b in order]
        vocab = symbols + [s + "</w>" for s in symbols] + ["".join(m) for m in merges] + [SOT, EOT]
        self.encoder = {tok: i for i, tok in enumerate(vocab)}
        self.decoder = {i: tok for tok, i in self.encoder.items()}
        self.bpe_ranks = {m: r for r, m in enumerate(merges)}
        self.cache = {SOT: SOT, EOT: EOT}

    def bpe(self, token):
        """Merge the symbols of one pre-token, lowest-ranked adjacent pair first, every occurrence of it left to right
        per round (simple_tokenizer.py:93-132).  Returns the pieces joined by spaces."""
        hit = self.cache.get(token)
        if hit is not None:
            return hit
        word = list(token[:-1]) + [token[-1] + "</w>"]
        ranks = self.bpe_ranks
        while len(word) > 1:
            best, best_rank = None, None
            for pair in zip(word, word[1:]):
                r = ranks.get(pair)
                if r is not None and (best_rank is None or r < best_rank):
                    best, best_rank = pair, r
            if best is None:
                break
            merged, i = [], 0
            while i < len(word):
                if i + 1 < len(word) and word[i] == best[0] and word[i + 1] == best[1]:
                    merged.append(best[0] + best[1])
                    i += 2
                else:
                    merged.append(word[i])
                    i += 1
            word = merged
        out = " ".join(word)
        self.cache[token] = out
        return out

    def encode(self, text):
        ids = []
        text = whitespace_clean(basic_clean(text)).lower()
        for tok in self.PATTERN.findall(text):
            sym = "".join(self.byte_encoder[b] for b in tok.encode("utf-8"))
            ids.extend(self.encoder[piece] for piece in self.bpe(sym).split(" "))
        return ids

    def decode(self, tokens):
        text = "".join(self.decoder[int(t)] for t in tokens)
        return bytearray(self.byte_decoder[c] for c in text).decode("utf-8", errors="replace").replace("</w>", " ")


_default = {}


def _default_tokenizer(name, cls):
    if name not in _default:
        _default[name] = cls()
    return _default[name]


def tokenize(texts, context_length=77, truncate=False, tokenizer=None):
    """simple_tokenizer.py:151-169: int64 [n, context_length], zero padded; too long raises unless truncate."""
    if isinstance(texts, str):
        texts = [texts]
    tk = tokenizer or _default_tokenizer("bpe", SimpleTokenizer)
    sot, eot = tk.encoder[SOT], tk.encoder[EOT]
    out = np.zeros((len(texts), context_length), dtype=np.int64)
    for i, text in enumerate(texts):
        ids = [sot] + tk.encode(text) + [eot]
        if len(ids) > context_length:
            if not truncate:
                raise RuntimeError("Input %s is too long for context length %d" % (text, context_length))
            ids = ids[:context_length]
            ids[-1] = eot
        out[i, :len(ids)] = ids
    return out


# ----------------------------------------------------------------------------------------------------------
# WordPiece (wukong-huahua, Chinese)
# ----------------------------------------------------------------------------------------------------------
def _is_cjk(cp):
    """clip_zh/utils.py:62-85: the CJK ideograph blocks."""
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or 0x2A700 <= cp <= 0x2B73F
            or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


def _is_punct(ch):
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:   # every non-alphanumeric ASCII symbol
        return True
    return unicodedata.category(ch).startswith("P")


def _is_space(ch):
    return ch in " \t\n\r" or unicodedata.category(ch) == "Zs"


def _is_control(ch):
    return ch not in "\t\n\r" and unicodedata.category(ch).startswith("C")


class WordpieceTokenizer(object):
    """BERT tokenisation as wukong-huahua's text tower uses it; specials [CLS] / [SEP] / [UNK]."""

    def __init__(self, vocab_path=None, sot_text="[CLS]", eot_text="[SEP]"):
        vocab_path = vocab_path or find_vocab("vocab_zh.txt")
        with open(vocab_path, encoding="utf-8") as f:
            vocab = [line.strip() for line in f]
        self.encoder = {tok: i for i, tok in enumerate(vocab)}
        self.decoder = {i: tok for tok, i in self.encoder.items()}
        self.unk_token = "[UNK]"
        self.never_split = [self.unk_token, SOT, EOT]    # the module-level defaults at class-creation time
        self.max_input_chars_per_word = 100
        self.sot_text, self.eot_text = sot_text, eot_text

    def _basic(self, text):
        chars = []
        for ch in text:                                   # drop NUL / U+FFFD / controls, blank out whitespace
            if ord(ch) in (0, 0xFFFD) or _is_control(ch):
                continue
            if _is_space(ch):
                chars.append(" ")
            elif _is_cjk(ord(ch)):
                chars.append(" " + ch + " ")               # every ideograph becomes its own word
            else:
                chars.append(ch)
        words = []
        for tok in "".join(chars).split():
            if tok in self.never_split:
                words.append(tok)
                continue
            tok = "".join(c for c in unicodedata.normalize("NFD", tok.lower()) if unicodedata.category(c) != "Mn")
            cur = ""
            for ch in tok:                                 # punctuation marks stand alone
                if _is_punct(ch):
                    if cur:
                        words.append(cur)
                    words.append(ch)
                    cur = ""
                else:
                    cur += ch
            if cur:
                words.append(cur)
        return " ".join(words).split()

    def _wordpiece(self, word):
        if len(word) > self.max_input_chars_per_word:
            return [self.unk_token]
        pieces, start = [], 0
        while start < len(word):
            end = len(word)
            while end > start:
                cand = ("##" if start else "") + word[start:end]
                if cand in self.encoder:
                    break
                end -= 1
            if end == start:
                return [self.unk_token]
            pieces.append(cand)
            start = end
        return pieces

    def text_tokenize(self, text):
        out = []
        for word in self._basic(text):
            out.extend([word] if word in self.never_split else self._wordpiece(word))
        return out

    def encode(self, text):
        unk = self.encoder.get(self.unk_token, 0)
        return [self.encoder.get(tok, unk) for tok in self.text_tokenize(text)]

    def decode(self, tokens):
        segs = [self.decoder.get(int(t), self.unk_token) for t in tokens]
        return "".join(s if s in self.never_split else s.lstrip("#") for s in segs)


def tokenize_zh(texts, tokenizer=None, context_length=77):
    """FrozenCLIPEmbedder_ZH.tokenize (wukong-huahua encoders/modules.py:31-50): [CLS] ids [SEP], zero padded; a prompt
    that does not fit keeps its first context_length - 1 ids and ends in [SEP]."""
    if isinstance(texts, str):
        texts = [texts]
    tk = tokenizer or _default_tokenizer("zh", WordpieceTokenizer)
    sot, eot = tk.encoder[tk.sot_text], tk.encoder[tk.eot_text]
    out = np.zeros((len(texts), context_length), dtype=np.int64)
    for i, text in enumerate(texts):
        ids = [sot] + tk.encode(text) + [eot]
        if len(ids) > context_length:
            ids = ids[:context_length - 1] + [eot]
        out[i, :len(ids)] = ids
    return out


# ----------------------------------------------------------------------------------------------------------
# Taichu-GLIDE: SentencePiece + padding / mask helpers
# ----------------------------------------------------------------------------------------------------------
class SentencePieceEncoder(object):
    """tokenizer/chinese_tokenizer.py:95-143 (`Encoder_SP`).  The model file (`cog-pretrain.model`, src/txt2img.py:89)
    is distributed with the GLIDE checkpoints, not with the reference repository."""

    def __init__(self, model_path):
        import sentencepiece as spm
        self.sp = spm.SentencePieceProcessor()
        if not self.sp.Load(model_path):
            raise FileNotFoundError(model_path)
        self.num_tokens = self.sp.vocab_size()

    @property
    def n_vocab(self):
        return self.num_tokens + 1           # + the end token

    @property
    def end_token(self):
        return self.n_vocab - 1

    def padded_tokens_and_mask(self, tokens, text_ctx):
        tokens = list(tokens[:text_ctx])
        pad = text_ctx - len(tokens)
        return tokens + [self.end_token] * pad, [True] * len(tokens) + [False] * pad

    def encode(self, text):
        return self.sp.EncodeAsIds(text)

    def decode(self, tokens):
        return self.sp.DecodeIds([int(t) for t in tokens if int(t) != self.end_token])

    def tokenize(self, text):
        return self.sp.EncodeAsPieces(text)

    def convert_tokens_to_ids(self, tokens):
        return [self.sp.PieceToId(t) for t in tokens]


def encode_and_pad(input_line, text_ctx, tokenizer):
    """caption_to_tokens.py:37-41."""
    return tokenizer.padded_tokens_and_mask(tokenizer.encode(input_line), text_ctx)


def convert_input_to_token_gen(input_line, pics_generated, text_ctx, tokenizer):
    """caption_to_tokens.py:21-27: [prompt] * n + [empty prompt] * n as int32 arrays (tokens, mask)."""
    tokens, mask = encode_and_pad(input_line, text_ctx, tokenizer)
    u_tokens, u_mask = tokenizer.padded_tokens_and_mask([], text_ctx)
    return (np.asarray([tokens] * pics_generated + [u_tokens] * pics_generated, dtype=np.int32),
            np.asarray([mask] * pics_generated + [u_mask] * pics_generated, dtype=np.int32))


def convert_input_to_token_super_res(input_line, pics_generated, text_ctx, tokenizer):
    """caption_to_tokens.py:30-34."""
    tokens, mask = encode_and_pad(input_line, text_ctx, tokenizer)
    return (np.asarray([tokens] * pics_generated, dtype=np.int32), np.asarray([mask] * pics_generated, dtype=np.int32))
