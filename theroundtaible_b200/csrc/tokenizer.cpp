// Byte-level tokenizer + fixed chat wrapper, shared with the oracle (oracle/tokenizer.py).
// The reference never tokenizes: it sends messages:[{role:"user",content}] and the external server
// applies its own template (/root/reference/src/adapters/local-llm.ts:162-165).  ids 0..4 = PAD, BOS,
// EOS, USER, ASSISTANT; id 5+b = byte b.  chat(prompt) = [BOS, USER] + bytes + [ASSISTANT].
#include <cstdio>
#include <string>
#include <vector>

static const int TOK_BOS = 1, TOK_USER = 3, TOK_ASSISTANT = 4, TOK_BYTE0 = 5;

int rt_tok_encode_chat(const char* s, size_t len, std::vector<int>& ids) {
    ids.clear();
    ids.reserve(len + 3);
    ids.push_back(TOK_BOS);
    ids.push_back(TOK_USER);
    for (size_t i = 0; i < len; ++i) ids.push_back(TOK_BYTE0 + (unsigned char)s[i]);
    ids.push_back(TOK_ASSISTANT);
    return (int)ids.size();
}

// length of a well-formed UTF-8 sequence starting at b[0] (0 if ill-formed), per Unicode Table 3-7
static int utf8_valid_len(const unsigned char* b, size_t avail) {
    const unsigned char c = b[0];
    if (c < 0x80) return 1;
    auto cont = [&](size_t i, unsigned lo, unsigned hi) { return i < avail && b[i] >= lo && b[i] <= hi; };
    if (c >= 0xC2 && c <= 0xDF) return cont(1, 0x80, 0xBF) ? 2 : 0;
    if (c == 0xE0) return cont(1, 0xA0, 0xBF) && cont(2, 0x80, 0xBF) ? 3 : 0;
    if ((c >= 0xE1 && c <= 0xEC) || c == 0xEE || c == 0xEF) return cont(1, 0x80, 0xBF) && cont(2, 0x80, 0xBF) ? 3 : 0;
    if (c == 0xED) return cont(1, 0x80, 0x9F) && cont(2, 0x80, 0xBF) ? 3 : 0;
    if (c == 0xF0) return cont(1, 0x90, 0xBF) && cont(2, 0x80, 0xBF) && cont(3, 0x80, 0xBF) ? 4 : 0;
    if (c >= 0xF1 && c <= 0xF3) return cont(1, 0x80, 0xBF) && cont(2, 0x80, 0xBF) && cont(3, 0x80, 0xBF) ? 4 : 0;
    if (c == 0xF4) return cont(1, 0x80, 0x8F) && cont(2, 0x80, 0xBF) && cont(3, 0x80, 0xBF) ? 4 : 0;
    return 0;
}

std::string rt_tok_decode(const int* ids, int n) {
    static const char* special[5] = {"<pad>", "<bos>", "<eos>", "<user>", "<assistant>"};
    std::string out;
    std::vector<unsigned char> run;
    auto flush = [&]() {
        size_t i = 0;
        while (i < run.size()) {
            const int l = utf8_valid_len(run.data() + i, run.size() - i);
            if (l > 0) {
                out.append((const char*)run.data() + i, l);
                i += l;
            } else {
                char buf[8];
                snprintf(buf, sizeof buf, "<0x%02X>", run[i]);
                out += buf;
                i += 1;
            }
        }
        run.clear();
    };
    for (int k = 0; k < n; ++k) {
        const int t = ids[k];
        if (t >= TOK_BYTE0 && t < TOK_BYTE0 + 256) {
            run.push_back((unsigned char)(t - TOK_BYTE0));
        } else {
            flush();
            if (t >= 0 && t < 5) {
                out += special[t];
            } else {
                char buf[24];
                snprintf(buf, sizeof buf, "<t%d>", t);
                out += buf;
            }
        }
    }
    flush();
    return out;
}
