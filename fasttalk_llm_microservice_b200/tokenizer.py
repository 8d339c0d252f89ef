"""Tokenizer + Llama-3 chat template + streaming detokeniser for the handler.

With Ollama / vLLM the reference never sees tokens (`app/core/vllm_handler.py:245-260` posts
`messages`; the engine container tokenises).  An in-process engine has to do it here.

* `HFTokenizer` wraps a real `tokenizer.json` (path in $FASTTALK_TOKENIZER_JSON) through the
  `tokenizers` package.
* `ByteTokenizer` is the fallback when no vocabulary file is on the box (there is none here:
  SURVEY.md Appendix D).  It is a bijection between text and ids: UTF-8 bytes are ids 0..255,
  every other non-special id maps to one private-use code point, so text the model produced
  re-encodes to exactly the tokens it was generated from (the re-sent history then matches
  the session's cached KV prefix) and one token is one non-empty chunk.
"""
from __future__ import annotations

import codecs
import os
import re
from typing import Dict, Iterable, List, Optional, Sequence

# Llama-3 special ids (public)
BOS, EOT_TEXT, START_HEADER, END_HEADER, EOT = 128000, 128001, 128006, 128007, 128009


_PRIVATE_USE = re.compile("([\U000F0000-\U0010FFFF])")


class ByteTokenizer:
    def __init__(self, vocab: int = 128256):
        self.vocab = vocab
        if vocab >= 128256:
            self.bos, self.end_of_text, self.start_header, self.end_header, self.eot = BOS, EOT_TEXT, START_HEADER, END_HEADER, EOT
        else:                       # tiny test vocabularies: specials at the top of the range
            assert vocab >= 512, "ByteTokenizer needs at least 512 ids"
            self.bos, self.end_of_text, self.start_header, self.end_header, self.eot = (vocab - 5, vocab - 4, vocab - 3, vocab - 2, vocab - 1)
        self.specials = {self.bos, self.end_of_text, self.start_header, self.end_header, self.eot}
        self.eos_ids = (self.eot, self.end_of_text)

    # ---- id <-> private-use code point for ids >= 256
    @staticmethod
    def _id_to_char(t: int) -> str:
        k = t - 256
        return chr(0xF0000 + k) if k < 0xFFFE else chr(0x100000 + (k - 0xFFFE))

    @staticmethod
    def _char_to_id(ch: str) -> Optional[int]:
        cp = ord(ch)
        if 0xF0000 <= cp < 0xF0000 + 0xFFFE:
            return cp - 0xF0000 + 256
        if 0x100000 <= cp < 0x100000 + 0xFFFE:
            return cp - 0x100000 + 0xFFFE + 256
        return None

    def encode(self, text: str) -> List[int]:
        # Runs of ordinary text are their UTF-8 bytes (one C-level call); only private-use code points, which
        # stand for ids >= 256, take the per-character path.  A 2k-character history used to cost ~1 ms of
        # Python per request here — 64 sessions arriving together spent more time in this loop than in prefill.
        if not text:
            return []
        if max(text) < "\U000F0000":
            return list(text.encode("utf-8"))
        out: List[int] = []
        for part in _PRIVATE_USE.split(text):
            if not part:
                continue
            t = self._char_to_id(part) if len(part) == 1 and ord(part) >= 0xF0000 else None
            if t is not None and t < self.vocab and t not in self.specials:
                out.append(t)
            else:
                out.extend(part.encode("utf-8"))
        return out

    def decode(self, ids: Iterable[int]) -> str:
        dec = self.stream_decoder()
        return "".join(dec.push(t) for t in ids) + dec.flush()

    def stream_decoder(self) -> "ByteStreamDecoder":
        return ByteStreamDecoder(self)

    def apply_chat_template(self, messages: Sequence[Dict[str, str]], add_generation_prompt: bool = True) -> List[int]:
        ids = [self.bos]
        for m in messages:
            ids += [self.start_header] + self.encode(str(m.get("role", "user"))) + [self.end_header]
            ids += self.encode("\n\n" + str(m.get("content") or "")) + [self.eot]
        if add_generation_prompt:
            ids += [self.start_header] + self.encode("assistant") + [self.end_header] + self.encode("\n\n")
        return ids


class ByteStreamDecoder:
    """One call per token; returns the text that token completes ('' while a multi-byte
    UTF-8 sequence is still open or for special tokens)."""

    def __init__(self, tok: ByteTokenizer):
        self.tok = tok
        self.dec = codecs.getincrementaldecoder("utf-8")(errors="replace")

    def push(self, t: int) -> str:
        if t in self.tok.specials:
            return self.dec.decode(b"", final=False)
        if t < 256:
            return self.dec.decode(bytes([t]), final=False)
        return self.dec.decode(b"", final=False) + ByteTokenizer._id_to_char(t)

    def flush(self) -> str:
        return self.dec.decode(b"", final=True)


class HFTokenizer:
    """A real tokenizer.json (e.g. Llama-3's) through `tokenizers`."""
    releases_gil = True        # the Rust encoder runs without the GIL: long histories are tokenised off the event loop

    def __init__(self, path: str):
        from tokenizers import Tokenizer
        self.tk = Tokenizer.from_file(path)
        self.vocab = self.tk.get_vocab_size()
        tid = self.tk.token_to_id
        self.bos = tid("<|begin_of_text|>") or BOS
        self.end_of_text = tid("<|end_of_text|>") or EOT_TEXT
        self.start_header = tid("<|start_header_id|>") or START_HEADER
        self.end_header = tid("<|end_header_id|>") or END_HEADER
        self.eot = tid("<|eot_id|>") or EOT
        self.specials = {self.bos, self.end_of_text, self.start_header, self.end_header, self.eot}
        self.eos_ids = (self.eot, self.end_of_text)

    def encode(self, text: str) -> List[int]:
        return self.tk.encode(text, add_special_tokens=False).ids

    def decode(self, ids: Iterable[int]) -> str:
        return self.tk.decode([t for t in ids if t not in self.specials], skip_special_tokens=True)

    def stream_decoder(self) -> "HFStreamDecoder":
        return HFStreamDecoder(self)

    apply_chat_template = ByteTokenizer.apply_chat_template


class HFStreamDecoder:
    """Incremental detokeniser: O(window) per token instead of re-decoding the whole reply.

    Byte-level BPE pieces do not decode independently (a token may end inside a UTF-8 sequence, and
    a leading-space rule may depend on the neighbour), so — as the vLLM / TGI detokenisers do — a
    short window of already-emitted ids is decoded together with the new ones and only the text past
    the window's own decoding is emitted: `prefix` = start of the window, `read` = ids already turned
    into text.  Output is held back while the tail decodes to U+FFFD (an open UTF-8 sequence)."""

    def __init__(self, tok: HFTokenizer):
        self.tok, self.ids, self.prefix, self.read = tok, [], 0, 0

    def _emit(self, final: bool) -> str:
        dec = self.tok.tk.decode
        done = dec(self.ids[self.prefix:self.read], skip_special_tokens=True)
        text = dec(self.ids[self.prefix:], skip_special_tokens=True)
        if len(text) <= len(done) or (not final and text.endswith("\ufffd")):
            return ""
        self.prefix, self.read = self.read, len(self.ids)
        if self.prefix > 64:                      # keep the id list short: drop what lies before the window
            del self.ids[:self.prefix]
            self.read -= self.prefix
            self.prefix = 0
        return text[len(done):]

    def push(self, t: int) -> str:
        if t in self.tok.specials:
            return ""
        self.ids.append(t)
        return self._emit(False)

    def flush(self) -> str:
        return self._emit(True) if self.read < len(self.ids) else ""


def load_tokenizer(vocab: int, path: Optional[str] = None, require: bool = False):
    """`require`: a real checkpoint is being served — refuse to fall back to the byte tokenizer, whose ids
    mean nothing to trained weights (set FASTTALK_ALLOW_BYTE_TOKENIZER=1 for random-init checkpoints)."""
    path = path or os.environ.get("FASTTALK_TOKENIZER_JSON")
    if path and os.path.exists(path):
        return HFTokenizer(path)
    if require and os.environ.get("FASTTALK_ALLOW_BYTE_TOKENIZER", "0") != "1":
        raise FileNotFoundError(
            "a checkpoint was given but no tokenizer.json was found beside it (or in $FASTTALK_TOKENIZER_JSON); "
            "serving it with the byte tokenizer would produce garbage text — set FASTTALK_ALLOW_BYTE_TOKENIZER=1 "
            "to do that on purpose (random-init test checkpoints)")
    return ByteTokenizer(vocab)
