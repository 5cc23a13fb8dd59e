"""Synthetic tokenizer vocabularies for the host-tokenisation parity tests.

The real GPT-2 / BERT vocabulary files are not available offline, so this script TRAINS small ones with the `tokenizers`
library (the backend of transformers' GPT2Tokenizer / BertTokenizer) on text found in this repository plus a few lines
of accented Latin, Greek, Cyrillic, CJK, Hangul and emoji, and writes them in the reference's file formats:

    tests/golden/vocab/gpt2_vocab.json + gpt2_merges.txt   (byte-level BPE, GPT2Tokenizer(vocab_file, merges_file))
    tests/golden/vocab/bert_vocab.txt                       (WordPiece, BertTokenizer(vocab_file))

The oracle of the tests is transformers' own tokenizer classes constructed from these files.
    python -m oracle.make_vocab
"""
import glob
import os

from tokenizers import Tokenizer, models, normalizers, pre_tokenizers, trainers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "vocab")
EXTRA = [
    "Héllo Wörld! Çà et là, naïve façade; El Niño año. ÀÉÎÕÜ ß Ø æ Œ.",
    "Ελληνικά: Σίσυφος, ΟΔΟΣ, ὀδός και λόγος. ΣΟΦΟΣ σοφός.",
    "Русский текст: Привет, мир! Ёлка и ёж. ДОМ дом.",
    "中文分词测试：分布式系统与共识算法。日本語のテキスト、カタカナとひらがな。",
    "한국어 텍스트 처리 테스트입니다.",
    "emoji 😀 😂 🤖 👍🏽 ❤️ and symbols © ® ™ € £ ¥ § ¶ • … — – “quotes” ‘single’ «guillemets»",
    "numbers 0123456789 ３４５ ²³ ½ Ⅻ ٠١٢٣ 12,345.67 3.14e-10 0x1F",
    "tabs\tand\nnewlines\r\nand   multiple    spaces  and trailing   ",
    "don't can't it's we're they've I'm you'll he'd DON'T CAN'T",
    "Answer the following question as a helpful tutor: what is raft consensus?",
]


def corpus():
    lines = list(EXTRA)
    for path in sorted(glob.glob(os.path.join(ROOT, "*.md"))) + sorted(glob.glob(os.path.join(ROOT, "oracle", "*.py"))):
        with open(path, encoding="utf-8", errors="ignore") as f:
            lines.extend(l.rstrip("\n") for l in f if l.strip())
    return lines


def main():
    os.makedirs(OUT, exist_ok=True)
    lines = corpus()
    bpe = Tokenizer(models.BPE())
    bpe.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    bpe.train_from_iterator(lines, trainers.BpeTrainer(vocab_size=3000, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                                       special_tokens=["<|endoftext|>"], show_progress=False))
    bpe.model.save(OUT, "gpt2")
    os.replace(os.path.join(OUT, "gpt2-vocab.json"), os.path.join(OUT, "gpt2_vocab.json"))
    os.replace(os.path.join(OUT, "gpt2-merges.txt"), os.path.join(OUT, "gpt2_merges.txt"))
    wp = Tokenizer(models.WordPiece(unk_token="[UNK]"))
    wp.normalizer = normalizers.BertNormalizer(clean_text=True, handle_chinese_chars=True, strip_accents=None, lowercase=True)
    wp.pre_tokenizer = pre_tokenizers.BertPreTokenizer()
    wp.train_from_iterator(lines, trainers.WordPieceTrainer(vocab_size=3000, show_progress=False,
                                                            special_tokens=["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"]))
    wp.model.save(OUT, "bert")
    os.replace(os.path.join(OUT, "bert-vocab.txt"), os.path.join(OUT, "bert_vocab.txt"))
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
