"""Byte-level tokenizer + fixed chat wrapper (TEST INFRASTRUCTURE; see oracle/__init__.py).

No tokenizer file exists on the box and the reference never tokenizes: it sends
``messages:[{role:"user",content:prompt}]`` (/root/reference/src/adapters/local-llm.ts:162-165)
and the external server applies its own template.  Oracle and engine share this definition
(theroundtaible_b200/csrc/tokenizer.cpp restates it):

  ids 0..4 : PAD, BOS, EOS, USER, ASSISTANT;  id 5+b : UTF-8 byte b (b in 0..255)
  chat(prompt) = [BOS, USER] + bytes(prompt) + [ASSISTANT]

Detokenising: byte tokens that form valid UTF-8 pass through; an invalid byte becomes
"<0xNN>"; any other id (a random-init model emits ids up to vocab-1) becomes "<tN>", so the
text is never empty -- the adapter rejects empty content (local-llm.ts:189-191).
"""
from __future__ import annotations

PAD, BOS, EOS, USER, ASSISTANT = 0, 1, 2, 3, 4
BYTE0 = 5


def encode_chat(prompt: str) -> list[int]:
    return [BOS, USER] + [BYTE0 + b for b in prompt.encode("utf-8")] + [ASSISTANT]


def _utf8_len(b0: int) -> int:
    if b0 < 0x80:
        return 1
    if 0xC2 <= b0 <= 0xDF:
        return 2
    if 0xE0 <= b0 <= 0xEF:
        return 3
    if 0xF0 <= b0 <= 0xF4:
        return 4
    return 0


def _valid_seq(bs: list[int]) -> bool:
    try:
        bytes(bs).decode("utf-8")
        return True
    except UnicodeDecodeError:
        return False


def decode(ids) -> str:
    out: list[str] = []
    run: list[int] = []

    def flush():
        i = 0
        while i < len(run):
            n = _utf8_len(run[i])
            if n and i + n <= len(run) and _valid_seq(run[i:i + n]):
                out.append(bytes(run[i:i + n]).decode("utf-8"))
                i += n
            else:
                out.append(f"<0x{run[i]:02X}>")
                i += 1
        run.clear()

    for t in ids:
        if BYTE0 <= t < BYTE0 + 256:
            run.append(t - BYTE0)
        else:
            flush()
            if t > ASSISTANT:
                out.append(f"<t{t}>")
            else:
                out.append(("<pad>", "<bos>", "<eos>", "<user>", "<assistant>")[t])
    flush()
    return "".join(out)
