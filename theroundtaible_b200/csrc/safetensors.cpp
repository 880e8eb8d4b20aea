// See safetensors.h.  A small recursive-descent JSON reader (the header is a flat object of objects) and an
// mmap of each file; nothing here touches the GPU.
#include "safetensors.h"

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>

#include <dirent.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace st {
namespace {

struct Json {
    const char* p;
    const char* end;
    const std::string& file;

    [[noreturn]] void fail(const char* what) const {
        throw std::runtime_error("safetensors header of '" + file + "' is malformed: " + what);
    }
    void ws() {
        while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
    }
    bool eat(char c) {
        ws();
        if (p < end && *p == c) { ++p; return true; }
        return false;
    }
    void expect(char c) {
        if (!eat(c)) fail("unexpected character");
    }
    std::string str() {
        ws();
        if (p >= end || *p != '"') fail("string expected");
        ++p;
        std::string out;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) fail("unterminated escape");
                switch (*p) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (end - p < 5) fail("short \\u escape");
                        unsigned cp = 0;
                        for (int i = 1; i <= 4; ++i) {
                            const char h = p[i];
                            cp = cp * 16 + (h >= '0' && h <= '9' ? h - '0' : h >= 'a' && h <= 'f' ? h - 'a' + 10 : h >= 'A' && h <= 'F' ? h - 'A' + 10 : 0);
                        }
                        p += 4;
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: out += *p;
                }
                ++p;
            } else {
                out += *p++;
            }
        }
        if (p >= end) fail("unterminated string");
        ++p;
        return out;
    }
    long long integer() {
        ws();
        const char* s = p;
        if (p < end && (*p == '-' || *p == '+')) ++p;
        while (p < end && *p >= '0' && *p <= '9') ++p;
        if (p == s) fail("integer expected");
        return std::stoll(std::string(s, p));
    }
    // skip any value (used for __metadata__ and unknown keys)
    void skip() {
        ws();
        if (p >= end) fail("value expected");
        if (*p == '"') { str(); return; }
        if (*p == '{' || *p == '[') {
            const char close = *p == '{' ? '}' : ']';
            const bool obj = *p == '{';
            ++p;
            if (eat(close)) return;
            do {
                if (obj) { str(); expect(':'); }
                skip();
            } while (eat(','));
            expect(close);
            return;
        }
        while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;    // number / true / false / null
    }
};

bool ends_with(const std::string& s, const char* suf) {
    const size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

}  // namespace

void Files::add_file(const std::string& file) {
    const int fd = open(file.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("weights file '" + file + "' not found");
    struct stat sb{};
    if (fstat(fd, &sb) != 0 || sb.st_size < 8) {
        close(fd);
        throw std::runtime_error("weights file '" + file + "' is too short to be a safetensors file");
    }
    void* base = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (base == MAP_FAILED) throw std::runtime_error("cannot map weights file '" + file + "'");
    maps_.push_back({base, (size_t)sb.st_size});
    const unsigned char* b = static_cast<const unsigned char*>(base);
    uint64_t hlen = 0;
    memcpy(&hlen, b, 8);
    if (hlen > (uint64_t)sb.st_size - 8) throw std::runtime_error("safetensors header of '" + file + "' is malformed: length exceeds the file");
    const unsigned char* payload = b + 8 + hlen;
    const size_t payload_len = (size_t)sb.st_size - 8 - (size_t)hlen;

    Json j{reinterpret_cast<const char*>(b + 8), reinterpret_cast<const char*>(b + 8 + hlen), file};
    j.expect('{');
    if (j.eat('}')) return;
    do {
        const std::string name = j.str();
        j.expect(':');
        if (name == "__metadata__") { j.skip(); continue; }
        Tensor t;
        long long off0 = -1, off1 = -1;
        j.expect('{');
        do {
            const std::string key = j.str();
            j.expect(':');
            if (key == "dtype") {
                t.dtype = j.str();
            } else if (key == "shape") {
                j.expect('[');
                if (!j.eat(']')) {
                    do t.shape.push_back(j.integer()); while (j.eat(','));
                    j.expect(']');
                }
            } else if (key == "data_offsets") {
                j.expect('[');
                off0 = j.integer();
                j.expect(',');
                off1 = j.integer();
                j.expect(']');
            } else {
                j.skip();
            }
        } while (j.eat(','));
        j.expect('}');
        if (off0 < 0 || off1 < off0 || (size_t)off1 > payload_len)
            throw std::runtime_error("tensor '" + name + "' in '" + file + "' has data_offsets outside the file");
        t.data = payload + off0;
        t.nbytes = (size_t)(off1 - off0);
        if (t.dtype == "BF16" || t.dtype == "F16" || t.dtype == "F32") {
            if ((size_t)t.numel() * t.elem_size() != t.nbytes)
                throw std::runtime_error("tensor '" + name + "' in '" + file + "': shape and data_offsets disagree");
        }
        tensors_[name] = std::move(t);
    } while (j.eat(','));
    j.expect('}');
}

Files::Files(const std::string& path) {
    struct stat sb{};
    if (stat(path.c_str(), &sb) != 0) throw std::runtime_error("weights path '" + path + "' not found");
    if (S_ISDIR(sb.st_mode)) {
        std::vector<std::string> files;
        if (DIR* d = opendir(path.c_str())) {
            while (const dirent* de = readdir(d))
                if (ends_with(de->d_name, ".safetensors")) files.push_back(path + "/" + de->d_name);
            closedir(d);
        }
        if (files.empty()) throw std::runtime_error("weights path '" + path + "': no *.safetensors file found in the directory");
        std::sort(files.begin(), files.end());
        for (const auto& f : files) add_file(f);
    } else {
        add_file(path);
    }
}

Files::~Files() {
    for (auto& m : maps_) munmap(m.base, m.len);
}

const Tensor* Files::find(const std::string& name) const {
    const auto it = tensors_.find(name);
    return it == tensors_.end() ? nullptr : &it->second;
}

const Tensor& Files::get(const std::string& name, std::initializer_list<long long> shape) const {
    const Tensor* t = find(name);
    if (!t) throw std::runtime_error("tensor '" + name + "' not found in the checkpoint");
    if (t->dtype != "BF16" && t->dtype != "F16" && t->dtype != "F32")
        throw std::runtime_error("tensor '" + name + "' has dtype " + t->dtype + "; BF16, F16 or F32 expected");
    if (t->shape != std::vector<long long>(shape)) {
        std::string got, want;
        for (long long d : t->shape) got += (got.empty() ? "" : "x") + std::to_string(d);
        for (long long d : shape) want += (want.empty() ? "" : "x") + std::to_string(d);
        throw std::runtime_error("tensor '" + name + "' has shape " + got + ", the model configuration needs " + want);
    }
    return *t;
}

}  // namespace st
