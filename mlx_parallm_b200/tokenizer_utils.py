"""TokenizerWrapper / streaming detokenizer with the surface the reference takes from
mlx_lm.tokenizer_utils (imported at mlx_parallm/utils.py:24; used at utils.py:449-455, 1016 and by
server/main.py): ``_tokenizer``, ``detokenizer.{reset, add_token, finalize, last_segment, text, tokens}``,
``eos_token_id`` and attribute pass-through to the wrapped HF tokenizer."""
from __future__ import annotations

from pathlib import Path


class NaiveStreamingDetokenizer:
    """Re-decodes the running token list and emits the newly completed text (mlx_lm's naive strategy)."""

    def __init__(self, tokenizer):
        self._tokenizer = tokenizer
        self.reset()

    def reset(self):
        self.offset = 0
        self._tokens = []
        self._text = ""
        self._current_tokens = []
        self._current_text = ""

    @property
    def tokens(self):
        return self._tokens

    def add_token(self, token):
        self._current_tokens.append(token)
        self._tokens.append(token)

    def finalize(self):
        self._text += self._tokenizer.decode(self._current_tokens)
        self._current_tokens = []
        self._current_text = ""

    @property
    def text(self):
        if self._current_tokens:
            self._current_text = self._tokenizer.decode(self._current_tokens)
            # hold back while the tail is an incomplete UTF-8 sequence
            if self._current_text.endswith("�"):
                self._current_text = self._current_text[:-1]
        if self._current_text and self._current_text[-1] == "\n":
            self._text += self._current_text
            self._current_tokens.clear()
            self._current_text = ""
        return self._text + self._current_text

    @property
    def last_segment(self):
        text = self.text
        if text and text[-1] != "�":
            segment = text[self.offset:]
            self.offset = len(text)
            return segment
        return ""


class TokenizerWrapper:
    def __init__(self, tokenizer, detokenizer_class=NaiveStreamingDetokenizer, eos_token_ids=None):
        self._tokenizer = tokenizer
        self._detokenizer = detokenizer_class(tokenizer)
        self._eos_token_ids = set(eos_token_ids) if eos_token_ids is not None else {tokenizer.eos_token_id}

    @property
    def detokenizer(self):
        return self._detokenizer

    @property
    def eos_token_ids(self):
        return self._eos_token_ids

    def __getattr__(self, attr):
        if attr == "detokenizer":
            return self._detokenizer
        if attr.startswith("_"):
            return self.__getattribute__(attr)
        return getattr(self._tokenizer, attr)

    def __setattr__(self, attr, value):
        if attr in ("_tokenizer", "_detokenizer", "_eos_token_ids"):
            super().__setattr__(attr, value)
        else:
            setattr(self._tokenizer, attr, value)


def load_tokenizer(model_path, tokenizer_config_extra={}, eos_token_ids=None):
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained(str(Path(model_path)), **tokenizer_config_extra)
    return TokenizerWrapper(tok, eos_token_ids=eos_token_ids)
