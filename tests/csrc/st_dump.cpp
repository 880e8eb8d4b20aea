// Test helper: lists what theroundtaible_b200/csrc/safetensors.cpp reads from a checkpoint path, one line per
// requested tensor: "<name> <dtype> <d0>x<d1> <fnv1a-64 of the tensor bytes>", or "ERROR <message>".
#include <cstdio>
#include <cstdint>
#include <exception>
#include <string>

#include "safetensors.h"

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    try {
        st::Files files(argv[1]);
        std::printf("COUNT %zu\n", files.size());
        for (int i = 2; i < argc; ++i) {
            const st::Tensor* t = files.find(argv[i]);
            if (!t) { std::printf("%s MISSING\n", argv[i]); continue; }
            uint64_t h = 1469598103934665603ull;
            for (size_t b = 0; b < t->nbytes; ++b) { h ^= t->data[b]; h *= 1099511628211ull; }
            std::string shape;
            for (long long d : t->shape) shape += (shape.empty() ? "" : "x") + std::to_string(d);
            std::printf("%s %s %s %016llx\n", argv[i], t->dtype.c_str(), shape.c_str(), (unsigned long long)h);
        }
    } catch (const std::exception& ex) {
        std::printf("ERROR %s\n", ex.what());
        return 1;
    }
    return 0;
}
