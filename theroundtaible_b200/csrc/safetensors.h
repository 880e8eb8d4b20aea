// Reader for .safetensors checkpoints (Hugging Face tensor names), host side only.
// No counterpart in the reference: its local-llm adapter names a model (src/types.ts:29-34) and the
// external server behind `endpoint` owns the weight files.  SURVEY.md 8(b) asks rt_engine_create for a
// "weight source (safetensors path or seeded generator)"; this is the first of the two.
//
// File format (published, safetensors 0.4): u64 little-endian header length N, N bytes of JSON
//   {"<name>": {"dtype": "BF16"|"F16"|"F32", "shape": [..], "data_offsets": [begin, end]}, ..., "__metadata__": {..}}
// then the tensor bytes, offsets relative to the end of the header, row-major.
#pragma once
#include <cstddef>
#include <map>
#include <string>
#include <vector>

namespace st {

struct Tensor {
    std::string dtype;               // "BF16", "F16", "F32"
    std::vector<long long> shape;
    const unsigned char* data = nullptr;   // into the file mapping; valid while the Files object lives
    size_t nbytes = 0;
    long long numel() const { long long n = 1; for (long long d : shape) n *= d; return n; }
    int elem_size() const { return dtype == "F32" ? 4 : 2; }
};

// One .safetensors file, or every *.safetensors file of a directory (sharded checkpoints), memory-mapped read-only.
// Throws std::runtime_error with a message that names the file / tensor.
class Files {
  public:
    explicit Files(const std::string& path);
    ~Files();
    Files(const Files&) = delete;
    Files& operator=(const Files&) = delete;
    const Tensor* find(const std::string& name) const;
    // find + dtype/shape check; throws when missing or mismatching ("tensor '<name>' ...")
    const Tensor& get(const std::string& name, std::initializer_list<long long> shape) const;
    size_t size() const { return tensors_.size(); }

  private:
    struct Mapping { void* base; size_t len; };
    void add_file(const std::string& file);
    std::vector<Mapping> maps_;
    std::map<std::string, Tensor> tensors_;
};

}  // namespace st
