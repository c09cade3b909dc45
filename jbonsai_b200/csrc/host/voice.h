// Host-side voice model: in-memory form of one `.htsvoice` file.
//
// Mirrors the data the reference keeps in `src/model/voice/{mod,model,tree,question,window}.rs`
// (Voice / Model / Tree / Question / Windows) — host glue, not kernel scope.
// Questions are matched as HTS wildcard patterns (`*`, `?`) against the full-context label
// string, which is what jlabel-question implements for well-formed labels.
#pragma once

#include <cstddef>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

namespace jb {

// Error codes cross the C ABI (include/jbonsai_b200.h); keep in sync.
enum ErrorCode : int {
  kOk = 0,
  kErrEmptyVoice = -1,      // ModelError::EmptyVoice        (model/mod.rs:31-46)
  kErrMetadata = -2,        // ModelError::MetadataError
  kErrIo = -3,              // ModelError::Io
  kErrParse = -4,           // ModelError::ParserError
  kErrOption = -5,          // EngineError::ParseOptionError (engine.rs:16-28)
  kErrLabelParse = -6,      // LabelError::JLabelParse       (label.rs:7-19)
  kErrLabelMissing = -7,    // LabelError::MissingLabel
  kErrLabelFloat = -8,      // LabelError::FloatParse
  kErrLabelLength = -9,     // LabelError::LengthMismatch
  kErrWeight = -10,         // WeightError::{InvalidSum,InvalidLength}
  kErrIndex = -11,          // tree/pdf index not found (reference: todo!() panic)
  kErrArgument = -12,       // bad argument at the C ABI (reference: panic!)
  kErrCuda = -13,           // CUDA runtime failure / no device (no CPU fallback)
  kErrUnsupported = -14,    // feature outside the built scope (e.g. stage != 0)
};

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& msg) : std::runtime_error(msg), code(c) {}
};

struct MeanVari {
  double mean;
  double vari;
};

// One leaf pdf: `parameters[k]` = (mean_k, variance_k), optional MSD weight.
struct ModelParameter {
  std::vector<MeanVari> parameters;
  bool has_msd = false;
  double msd = 0.0;
};

// Pre-compiled HTS wildcard pattern.
struct Pattern {
  std::string text;
  // Segments between '*'; a segment may contain '?'.
  std::vector<std::string> segs;
  bool anchored_front = false;  // pattern does not start with '*'
  bool anchored_back = false;   // pattern does not end with '*'
  bool has_qmark = false;
  explicit Pattern(const std::string& p);
  bool match(const char* s, size_t n) const;
};

struct Question {
  std::vector<Pattern> patterns;
  bool test(const std::string& label) const {
    for (const auto& p : patterns)
      if (p.match(label.data(), label.size())) return true;
    return false;
  }
};

struct TreeNode {
  bool leaf = false;
  int question = -1;     // index into Model::questions
  size_t yes = 0, no = 0;
  size_t pdf_index = 0;  // 1-based
};

struct Tree {
  size_t state = 0;  // HTS state number, 2-based
  std::vector<TreeNode> nodes;
};

// Leaf indices already resolved for a label string, per tree (host front-end cache: a decision
// tree walk costs a few hundred wildcard matches, and batches repeat labels). Bounded; sharded
// so that the front-end threads of a batch do not serialise on it. Results are exactly the tree
// walk's (the Model computes on a miss). Keyed by the 64-bit hash of the label, which the caller
// computes ONCE per label and hands to every model it asks (a 200-byte label hashed 32 times was
// 40 % of the index-mode front-end); the entry keeps the label, and a second label with the same
// hash is simply not cached.
struct LabelCache {
  static constexpr size_t kShards = 64;
  static constexpr size_t kMaxPerShard = 4096;
  struct Entry {
    std::string label;
    std::vector<int32_t> pdf;  // per tree: 0 = not resolved yet, -1 = no leaf, else the 1-based pdf index
  };
  struct Identity {
    size_t operator()(uint64_t h) const { return (size_t)h; }
  };
  struct Shard {
    std::mutex mu;
    std::unordered_map<uint64_t, Entry, Identity> map;
  };
  Shard shards[kShards];
};
inline uint64_t label_hash(const std::string& label) { return (uint64_t)std::hash<std::string>()(label); }

struct Model {
  std::vector<Question> questions;
  std::vector<Tree> trees;
  std::vector<std::vector<ModelParameter>> pdf;  // [tree][pdf_index-1]
  std::shared_ptr<LabelCache> cache = std::make_shared<LabelCache>();

  // (tree_index + 2, pdf_index) like Model::get_index (voice/model.rs:50-67); -1 = None.
  void get_index(size_t state_index, const std::string& label, long* tree_index,
                 long* pdf_index) const;
  // The pdf index of `label` in EVERY tree of the model (pdf_index[t] for trees[t]; -1 = no leaf), one cache
  // lookup: what five get_index calls (one per state) return. `hash` = label_hash(label).
  void get_all(const std::string& label, uint64_t hash, int32_t* pdf_index) const;
  const ModelParameter& get_parameter(size_t state_index, const std::string& label) const;

 private:
  long walk(const Tree& tree, const std::string& label) const;
};

struct Window {
  std::vector<double> coefficients;
  size_t width() const { return coefficients.size(); }
  size_t left_width() const { return width() / 2; }
  size_t right_width() const { return width() - left_width() - 1; }
};

struct Windows {
  std::vector<Window> windows;
  size_t size() const { return windows.size(); }
  size_t max_width() const {
    size_t m = 0;
    for (const auto& w : windows) m = w.width() > m ? w.width() : m;
    return m / 2;
  }
};

struct GlobalModelMetadata {
  std::string hts_voice_version;
  size_t sampling_frequency = 0;
  size_t frame_period = 0;
  size_t num_states = 0;
  size_t num_streams = 0;
  std::vector<std::string> stream_type;
  std::string fullcontext_format;
  std::string fullcontext_version;
  std::vector<std::string> gv_off_context_patterns;
  Question gv_off_context;
  bool operator==(const GlobalModelMetadata& o) const;
};

struct StreamModelMetadata {
  size_t vector_length = 0;
  size_t num_windows = 0;
  bool is_msd = false;
  bool use_gv = false;
  std::vector<std::string> option;
  bool operator==(const StreamModelMetadata& o) const {
    return vector_length == o.vector_length && num_windows == o.num_windows &&
           is_msd == o.is_msd && use_gv == o.use_gv && option == o.option;
  }
};

struct StreamModels {
  StreamModelMetadata metadata;
  Model stream_model;
  bool has_gv = false;
  Model gv_model;
  Windows windows;
};

struct Voice {
  GlobalModelMetadata metadata;
  Model duration_model;
  std::vector<StreamModels> stream_models;
};

// `.htsvoice` parser (reference: model/parser/mod.rs:58-174). Throws jb::Error(kErrParse).
std::shared_ptr<Voice> parse_htsvoice(const uint8_t* bytes, size_t len);
std::vector<uint8_t> read_file(const std::string& path);  // throws kErrIo

}  // namespace jb
