"""CPU: the C++ host tokenizers (csrc/tokenizer.cpp) against transformers' own GPT2Tokenizer / BertTokenizer built from
the same vocabulary files (tests/golden/vocab/, trained by oracle/make_vocab.py): ids bit-identical, decoded text equal,
for curated edge cases and seeded random text over many scripts.  """
import os
import random
import time

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VOCAB = os.path.join(ROOT, "tests", "golden", "vocab")
PROMPT = "Answer the following question as a helpful tutor: {query}"  # REF/tutoring_server.py:18

CURATED = [
    "", " ", "  ", "a", " a", "a ", "  a  b  ", "\t", "\n\n", " \t \n x", "x \t\n", "\r\n",
    "Hello world", " Hello  world!!  ", "don't can't it's we're they've I'm you'll he'd", "DON'T 'S 'T 'Re 'VE",
    "'s't're've'm'll'd'", "''''", "a'b", "123abc456", " 12 34", "3.14e-10 0x1F 1,000,000",
    "²³½Ⅻ٠١٢",
    "Héllo Wörld! Çà et là, naïve façade",
    "ÀÉÎÕÜ ß Ø æ Œ ĳ ŉ",
    "İstanbul ıi İİ",
    "ΣΟΦΟΣ σοφός ΟΔΟΣ ὀδός ς Σ",
    "Привет, мир! Ёлка ёж ДОМ",
    "中文分词测试：分布式系统",
    "日本語のテキスト、カタカナ",
    "한국어 텍스트 처리", "mixed中文and英文123",
    "emoji \U0001f600\U0001f602\U0001f916\U0001f44d\U0001f3fd❤️ \U0001f1fa\U0001f1f8 \U0001f468‍\U0001f469‍\U0001f467",
    "© ® ™ € £ ¥ § ¶ • … — – “q” ‘s’ «g»",
    "a b c　d e fg",
    "zero​width‍join﻿bom", "ctrl\x00\x01\x1f\x7f\x9fend", "soft­hyphen", "� replacement",
    "combining á ë ñ ộ", "x" * 150, "word " * 300, "sup" + "er" * 60 + "califragilistic",
    "<|endoftext|>", "a<|endoftext|>b", "<|endoftext|><|endoftext|>", "< |endoftext|>", "[CLS] hi [SEP] [MASK] [PAD] [UNK]",
    "[SEP]", "x[SEP]y", "[sep] [Sep]", "[[SEP]]", "tabs\tand\nnewlines\r\nand   multiple    spaces  and trailing   ",
    PROMPT.format(query="what is raft consensus and how does leader election work?"),
    "How do I compute the gradient of a softmax cross-entropy loss w.r.t. the logits?  (show steps)",
]

ALPHABETS = [
    "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ", "0123456789", " \t\n", "    ", ".,;:!?'\"()[]{}-_/\\@#$%^&*+=<>|~`",
    "àáâãäåçèéêëìíîïñòóôõöùúûüýÿÀÉÎÕÜßØæŒ",
    "αβγδεζηθικλμνξοπρσςτυφχψωΑΒΓΔΣΩ",
    "абвгдеёжзийклмнопрстуфхцчшщъыьэюяАБВГДЁЖ",
    "中文分词测试分布式系统与共识算法日本語", "あいうえおカタカナ",
    "한국어텍스트처리", "\U0001f600\U0001f602\U0001f916\U0001f44d❤\U0001f389\U0001f680",
    "©®™€£¥§¶•…—–“”‘’«»",
    "  　​­̣́̈", "'s't're've'm'll'd",
]


def random_text(rng, n):
    out = []
    for _ in range(n):
        alpha = rng.choice(ALPHABETS) if rng.random() < 0.5 else ALPHABETS[0]
        out.append("".join(rng.choice(alpha) for _ in range(rng.randint(1, 9))))
        if rng.random() < 0.6:
            out.append(" ")
    return "".join(out)


@pytest.fixture(scope="module")
def gpt2_pair(pkg):
    from transformers import GPT2Tokenizer
    v, m = os.path.join(VOCAB, "gpt2_vocab.json"), os.path.join(VOCAB, "gpt2_merges.txt")
    return pkg.B200GPT2Tokenizer(v, m), GPT2Tokenizer(vocab=v, merges=m)


@pytest.fixture(scope="module")
def bert_pair(pkg):
    from transformers import BertTokenizer
    v = os.path.join(VOCAB, "bert_vocab.txt")
    vocab = {l.rstrip("\n"): i for i, l in enumerate(open(v, encoding="utf-8"))}
    return pkg.B200BertTokenizer(v), BertTokenizer(vocab=vocab)


def test_gpt2_bpe_ids_and_decode_match_transformers(gpt2_pair):
    mine, hf = gpt2_pair
    assert len(mine) == len(hf) and mine.eos_token_id == hf.eos_token_id
    rng = random.Random(1)
    texts = CURATED + [random_text(rng, rng.randint(1, 40)) for _ in range(400)]
    for text in texts:
        ref = hf.encode(text)
        got = mine.encode(text)
        assert got == ref, (text, got[:20], ref[:20])
        assert mine.decode(got, skip_special_tokens=True) == hf.decode(ref, skip_special_tokens=True), text
        assert mine.decode(got) == hf.decode(ref), text
    # the reference's call shapes (REF/tutoring_server.py:20, 30)
    prompt = PROMPT.format(query="what is 2 + 2?")
    t = mine.encode(prompt, return_tensors="pt")
    assert t.dtype == torch.int64 and torch.equal(t, hf.encode(prompt, return_tensors="pt"))
    assert mine.decode(t[0], skip_special_tokens=True) == hf.decode(t[0], skip_special_tokens=True) == prompt
    # ids that do not form valid UTF-8 on their own decode with the replacement character, like transformers
    odd = hf.encode("é\U0001f600")[:-1]
    assert mine.decode(odd) == hf.decode(odd)


def test_bert_wordpiece_ids_match_transformers(bert_pair):
    mine, hf = bert_pair
    rng = random.Random(2)
    texts = CURATED + [random_text(rng, rng.randint(1, 60)) for _ in range(400)]
    for text in texts:
        ref = hf(text, return_tensors="pt", padding=True, truncation=True, max_length=512)
        got = mine(text, return_tensors="pt", padding=True, truncation=True, max_length=512)
        assert torch.equal(got["input_ids"], ref["input_ids"]), (text, got["input_ids"][0, :20], ref["input_ids"][0, :20])
        assert torch.equal(got["token_type_ids"], ref["token_type_ids"]) and torch.equal(got["attention_mask"], ref["attention_mask"])
    long_text = "consensus " * 700  # truncation=True, max_length=512 (REF/lms_server.py:98)
    ref = hf(long_text, return_tensors="pt", padding=True, truncation=True, max_length=512)["input_ids"]
    got = mine(long_text, return_tensors="pt", padding=True, truncation=True, max_length=512)["input_ids"]
    assert got.shape == (1, 512) and torch.equal(got, ref)
    assert mine("", return_tensors="pt")["input_ids"].tolist() == hf("", return_tensors="pt")["input_ids"].tolist()  # the empty query
    for ml in (2, 3, 17):
        assert mine("a b c d e f g h", max_length=ml)["input_ids"].tolist() == \
            hf("a b c d e f g h", return_tensors="pt", truncation=True, max_length=ml)["input_ids"].tolist()


def test_tokenizer_throughput(gpt2_pair, bert_pair):
    """us per request next to transformers' tokenizers (reported; the only assertion is that the C path is not slower
    than 3x the Rust one, i.e. it has not become the bottleneck it was built to remove)."""
    mine_g, hf_g = gpt2_pair
    mine_b, hf_b = bert_pair
    rng = random.Random(3)
    queries = [PROMPT.format(query=random_text(rng, 25)) for _ in range(200)]
    docs = [random_text(rng, 400) for _ in range(50)]

    def clock(fn, items):
        fn(items[0])
        t0 = time.perf_counter()
        for x in items:
            fn(x)
        return (time.perf_counter() - t0) / len(items) * 1e6
    enc = [mine_g.encode(q) for q in queries]
    rows = [("gpt2 encode (prompt + query)", clock(lambda q: mine_g.encode(q, return_tensors="pt"), queries),
             clock(lambda q: hf_g.encode(q, return_tensors="pt"), queries)),
            ("gpt2 decode", clock(lambda e: mine_g.decode(e, skip_special_tokens=True), enc),
             clock(lambda e: hf_g.decode(e, skip_special_tokens=True), enc)),
            ("bert encode (assignment text, 512 cap)", clock(lambda d: mine_b(d, return_tensors="pt"), docs),
             clock(lambda d: hf_b(d, return_tensors="pt", padding=True, truncation=True, max_length=512), docs))]
    for name, a, b in rows:
        print(f"{name}: C library {a:.1f} us / request, transformers {b:.1f} us / request")
        assert a < 3 * b + 50
