"""A tokenizer with the slice of the HuggingFace surface the reference's text front-ends touch (utils.py:473-543, 983-1075,
1087-1346): ``__call__`` with left/right padding + attention masks, ``encode / decode / batch_decode``,
``apply_chat_template``, ``eos/pad`` tokens, ``padding_side``, ``model_max_length``.

Vocabulary: token id i <-> the word "w<i>"; id ``eos_token_id`` <-> "<eos>".  Text is words separated by single spaces, so
``decode(encode(t)) == t`` and transcripts can be compared as strings or as ids."""
import numpy as np


class FakeTokenizer:
    def __init__(self, vocab_size: int, eos_token_id: int = 0, pad_token_id=None, model_max_length: int = 4096):
        self.vocab_size = vocab_size
        self.eos_token_id = eos_token_id
        self.eos_token = "<eos>"
        self.pad_token_id = pad_token_id
        self.pad_token = None if pad_token_id is None else self._word(pad_token_id)
        self.padding_side = "right"
        self.model_max_length = model_max_length
        self.chat_template_calls = 0
        self.encode_calls = 0

    # -- words
    def _word(self, i: int) -> str:
        return self.eos_token if i == self.eos_token_id else f"w{i}"

    def _id(self, w: str) -> int:
        return self.eos_token_id if w == self.eos_token else int(w[1:])

    def encode(self, text: str, **kw):
        self.encode_calls += 1
        return [self._id(w) for w in text.split()]

    def decode(self, ids, skip_special_tokens: bool = False, **kw):
        ids = [int(i) for i in ids]
        if skip_special_tokens:
            ids = [i for i in ids if i != self.eos_token_id]
        return " ".join(self._word(i) for i in ids)

    def batch_decode(self, rows, **kw):
        return [self.decode(r, **kw) for r in rows]

    def apply_chat_template(self, messages, add_generation_prompt: bool = True, tokenize: bool = False, **kw):
        """<user> = w1, <assistant> = w2: 'w1 <content> w2'."""
        self.chat_template_calls += 1
        parts = []
        for m in messages:
            parts += ["w1" if m["role"] == "user" else "w2", m["content"]]
        if add_generation_prompt:
            parts.append("w2")
        text = " ".join(parts)
        return self.encode(text) if tokenize else text

    def __call__(self, texts, padding=False, return_tensors=None, truncation=False, max_length=None, **kw):
        rows = [self.encode(t) for t in texts]
        if truncation and max_length is not None:
            rows = [r[:max_length] for r in rows]
        if padding:
            n = max(len(r) for r in rows)
            pad = self.pad_token_id if self.pad_token_id is not None else self.eos_token_id
            if self.padding_side == "left":
                ids = [[pad] * (n - len(r)) + r for r in rows]
                mask = [[0] * (n - len(r)) + [1] * len(r) for r in rows]
            else:
                ids = [r + [pad] * (n - len(r)) for r in rows]
                mask = [[1] * len(r) + [0] * (n - len(r)) for r in rows]
        else:
            ids, mask = rows, [[1] * len(r) for r in rows]
        if return_tensors == "np":
            return {"input_ids": np.array(ids, dtype=np.int64), "attention_mask": np.array(mask, dtype=np.int64)}
        return {"input_ids": ids, "attention_mask": mask}
