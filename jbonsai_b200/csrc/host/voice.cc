// `.htsvoice` reader, wildcard questions and decision-tree lookup (host glue).
// Behaviour follows the reference's `src/model/parser/**` and `src/model/voice/**`; the
// implementation is a hand-written scanner (the reference uses nom + serde).
#include "voice.h"

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>

namespace jb {

// ---------------------------------------------------------------- wildcard patterns
Pattern::Pattern(const std::string& p) : text(p) {
  anchored_front = p.empty() || p.front() != '*';
  anchored_back = p.empty() || p.back() != '*';
  std::string cur;
  bool any_star = false;
  for (char c : p) {
    if (c == '*') {
      any_star = true;
      if (!cur.empty()) segs.push_back(cur);
      cur.clear();
    } else {
      if (c == '?') has_qmark = true;
      cur.push_back(c);
    }
  }
  if (!cur.empty()) segs.push_back(cur);
  if (!any_star) anchored_front = anchored_back = true;
}

static inline bool seg_at(const std::string& seg, const char* s) {
  for (size_t i = 0; i < seg.size(); ++i)
    if (seg[i] != '?' && seg[i] != s[i]) return false;
  return true;
}

static inline long seg_find(const std::string& seg, bool qmark, const char* s, size_t from,
                            size_t n) {
  if (seg.size() > n) return -1;
  if (!qmark) {
    if (from + seg.size() > n) return -1;
    const void* r = memmem(s + from, n - from, seg.data(), seg.size());
    return r ? (long)((const char*)r - s) : -1;
  }
  for (size_t i = from; i + seg.size() <= n; ++i)
    if (seg_at(seg, s + i)) return (long)i;
  return -1;
}

bool Pattern::match(const char* s, size_t n) const {
  if (segs.empty()) {
    // "" matches only "", "*"/"**" match everything
    return (anchored_front && anchored_back) ? n == 0 : true;
  }
  size_t first = 0, last = segs.size();
  size_t lo = 0, hi = n;
  if (anchored_front) {
    const std::string& g = segs[0];
    if (g.size() > n || !seg_at(g, s)) return false;
    lo = g.size();
    first = 1;
    if (segs.size() == 1 && anchored_back) return g.size() == n;
  }
  if (anchored_back) {
    if (first >= last) return false;  // unreachable: handled above
    const std::string& g = segs[last - 1];
    if (g.size() > hi - lo) return false;
    if (!seg_at(g, s + hi - g.size())) return false;
    hi -= g.size();
    last -= 1;
  }
  for (size_t k = first; k < last; ++k) {
    long pos = seg_find(segs[k], has_qmark, s, lo, hi);
    if (pos < 0) return false;
    lo = (size_t)pos + segs[k].size();
  }
  return true;
}

// ---------------------------------------------------------------- tree lookup
long Model::walk(const Tree& tree, const std::string& label) const {
  size_t node = 0;
  while (node < tree.nodes.size()) {
    const TreeNode& nd = tree.nodes[node];
    if (nd.leaf) return (long)nd.pdf_index;
    node = questions[(size_t)nd.question].test(label) ? nd.yes : nd.no;
  }
  return -1;
}

void Model::get_index(size_t state_index, const std::string& label, long* tree_index,
                      long* pdf_index) const {
  long ti = -1;
  for (size_t i = 0; i < trees.size(); ++i)
    if (trees[i].state == state_index) {
      ti = (long)i;
      break;
    }
  const size_t tsel = ti < 0 ? 0 : (size_t)ti;
  const Tree& tree = trees[tsel];
  *tree_index = ti < 0 ? -1 : ti + 2;
  const uint64_t h = label_hash(label);
  LabelCache::Shard& sh = cache->shards[h % LabelCache::kShards];
  bool foreign = false;  // the slot of this hash belongs to another label: no caching
  {
    std::lock_guard<std::mutex> g(sh.mu);
    auto it = sh.map.find(h);
    if (it != sh.map.end()) {
      if (it->second.label != label) foreign = true;
      else if (it->second.pdf[tsel] != 0) {
        *pdf_index = it->second.pdf[tsel];
        return;
      }
    }
  }
  const long pi = walk(tree, label);
  *pdf_index = pi;
  if (foreign) return;
  std::lock_guard<std::mutex> g(sh.mu);
  auto it = sh.map.find(h);
  if (it == sh.map.end()) {
    if (sh.map.size() >= LabelCache::kMaxPerShard) return;
    it = sh.map.emplace(h, LabelCache::Entry{label, std::vector<int32_t>(trees.size(), 0)}).first;
  } else if (it->second.label != label) {
    return;
  }
  it->second.pdf[tsel] = (int32_t)pi;
}

void Model::get_all(const std::string& label, uint64_t h, int32_t* out) const {
  const size_t nt = trees.size();
  LabelCache::Shard& sh = cache->shards[h % LabelCache::kShards];
  bool have = false, foreign = false;
  {
    std::lock_guard<std::mutex> g(sh.mu);
    auto it = sh.map.find(h);
    if (it != sh.map.end()) {
      if (it->second.label != label) foreign = true;
      else {
        have = true;
        for (size_t t = 0; t < nt; ++t) out[t] = it->second.pdf[t];
      }
    }
  }
  bool complete = have;
  if (!have)
    for (size_t t = 0; t < nt; ++t) out[t] = 0;
  else
    for (size_t t = 0; t < nt; ++t) complete = complete && out[t] != 0;
  if (complete) return;
  for (size_t t = 0; t < nt; ++t)
    if (out[t] == 0) out[t] = (int32_t)walk(trees[t], label);
  if (foreign) return;
  std::lock_guard<std::mutex> g(sh.mu);
  auto it = sh.map.find(h);
  if (it == sh.map.end()) {
    if (sh.map.size() >= LabelCache::kMaxPerShard) return;
    it = sh.map.emplace(h, LabelCache::Entry{label, std::vector<int32_t>(nt, 0)}).first;
  } else if (it->second.label != label) {
    return;
  }
  for (size_t t = 0; t < nt; ++t) it->second.pdf[t] = out[t];
}

const ModelParameter& Model::get_parameter(size_t state_index, const std::string& label) const {
  long ti, pi;
  get_index(state_index, label, &ti, &pi);
  if (ti < 0 || pi < 1 || (size_t)(ti - 2) >= pdf.size() || (size_t)(pi - 1) >= pdf[ti - 2].size())
    throw Error(kErrIndex, "tree/pdf index not found for label: " + label);
  return pdf[(size_t)ti - 2][(size_t)pi - 1];
}

bool GlobalModelMetadata::operator==(const GlobalModelMetadata& o) const {
  return hts_voice_version == o.hts_voice_version &&
         sampling_frequency == o.sampling_frequency && frame_period == o.frame_period &&
         num_states == o.num_states && num_streams == o.num_streams &&
         stream_type == o.stream_type && fullcontext_format == o.fullcontext_format &&
         fullcontext_version == o.fullcontext_version &&
         gv_off_context_patterns == o.gv_off_context_patterns;
}

// ---------------------------------------------------------------- file reading
std::vector<uint8_t> read_file(const std::string& path) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) throw Error(kErrIo, "Io failed: cannot open " + path);
  std::vector<uint8_t> buf;
  uint8_t tmp[1 << 16];
  size_t n;
  while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
  fclose(f);
  return buf;
}

// ---------------------------------------------------------------- header sections
namespace {

[[noreturn]] void parse_fail(const std::string& what) {
  throw Error(kErrParse, "Parser returned error: " + what);
}

struct Header {
  // key -> values ; key[idx] -> values
  std::map<std::string, std::vector<std::string>> plain;
  std::map<std::string, std::map<std::string, std::vector<std::string>>> indexed;
};

// "v1,v2" with optional double quotes around items (header/de.rs:101-116).
std::vector<std::string> split_values(const std::string& v) {
  std::vector<std::string> out;
  size_t i = 0;
  if (v.empty()) return out;
  while (i <= v.size()) {
    std::string item;
    if (i < v.size() && v[i] == '"') {
      size_t e = v.find('"', i + 1);
      if (e == std::string::npos) parse_fail("unterminated string in header");
      item = v.substr(i + 1, e - i - 1);
      i = e + 1;
    } else {
      size_t e = v.find(',', i);
      if (e == std::string::npos) e = v.size();
      item = v.substr(i, e - i);
      i = e;
    }
    out.push_back(item);
    if (i >= v.size()) break;
    if (v[i] != ',') parse_fail("expected ',' in header value");
    ++i;
  }
  return out;
}

Header parse_header(const std::string& text) {
  Header h;
  size_t pos = 0;
  while (pos < text.size()) {
    size_t e = text.find('\n', pos);
    if (e == std::string::npos) e = text.size();
    std::string line = text.substr(pos, e - pos);
    pos = e + 1;
    if (line.empty()) continue;
    size_t colon = line.find(':');
    if (colon == std::string::npos) parse_fail("header line without ':' : " + line);
    std::string key = line.substr(0, colon), val = line.substr(colon + 1);
    size_t lb = key.find('[');
    if (lb != std::string::npos) {
      size_t rb = key.find(']', lb);
      if (rb == std::string::npos) parse_fail("bad header key " + key);
      h.indexed[key.substr(0, lb)][key.substr(lb + 1, rb - lb - 1)] = split_values(val);
    } else {
      h.plain[key] = split_values(val);
    }
  }
  return h;
}

const std::vector<std::string>& need(const Header& h, const std::string& key) {
  auto it = h.plain.find(key);
  if (it == h.plain.end()) parse_fail("missing field " + key);
  return it->second;
}

size_t to_usize(const std::string& s) {
  if (s.empty()) parse_fail("expected integer");
  for (char c : s)
    if (c < '0' || c > '9') parse_fail("expected integer, got " + s);
  return (size_t)strtoull(s.c_str(), nullptr, 10);
}

std::string one_string(const std::vector<std::string>& v) { return v.empty() ? "" : v[0]; }

bool to_bool(const std::vector<std::string>& v) {
  if (v.size() != 1 || (v[0] != "0" && v[0] != "1")) parse_fail("expected bool (0/1)");
  return v[0] == "1";
}

std::pair<size_t, size_t> to_range(const std::string& s) {
  size_t d = s.find('-');
  if (d == std::string::npos) parse_fail("expected a-b range, got " + s);
  return {to_usize(s.substr(0, d)), to_usize(s.substr(d + 1))};
}

// ---------------------------------------------------------------- tree / question text
struct Scanner {
  const char* p;
  const char* end;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  }
  bool starts(const char* lit) const {
    size_t n = strlen(lit);
    return (size_t)(end - p) >= n && memcmp(p, lit, n) == 0;
  }
  void expect(char c) {
    if (p >= end || *p != c) parse_fail(std::string("expected '") + c + "' in tree section");
    ++p;
  }
  std::string token() {  // run of non-space characters
    const char* s = p;
    while (p < end && *p != ' ' && *p != '\n' && *p != '\t' && *p != '\r') ++p;
    if (s == p) parse_fail("expected token in tree section");
    return std::string(s, p);
  }
};

struct RawChild {
  bool is_node;
  long id;  // node id (<= 0) or 1-based pdf index
};

RawChild parse_child(std::string tok) {
  if (tok.size() >= 2 && tok.front() == '"' && tok.back() == '"') tok = tok.substr(1, tok.size() - 2);
  // signed integer => node id (parser/model/tree.rs:44-55,77-78)
  {
    size_t i = 0;
    if (i < tok.size() && tok[i] == '-') ++i;
    bool digits = i < tok.size();
    for (size_t k = i; k < tok.size(); ++k)
      if (!isdigit((unsigned char)tok[k])) digits = false;
    if (digits) return {true, strtol(tok.c_str(), nullptr, 10)};
  }
  // otherwise the trailing digit run of the name is the pdf index (tree.rs:60-75)
  std::string id;
  for (char c : tok) {
    if (isdigit((unsigned char)c))
      id.push_back(c);
    else
      id.clear();
  }
  if (id.empty()) parse_fail("leaf name without trailing pdf index: " + tok);
  return {false, strtol(id.c_str(), nullptr, 10)};
}

struct RawNode {
  long id;
  std::string question;
  RawChild yes, no;
};

void parse_tree_section(const uint8_t* data, size_t len, Model* model) {
  Scanner sc{(const char*)data, (const char*)data + len};
  std::map<std::string, int> qindex;
  sc.ws();
  while (sc.p < sc.end && sc.starts("QS")) {
    sc.p += 2;
    sc.ws();
    std::string name = sc.token();
    sc.ws();
    sc.expect('{');
    Question q;
    for (;;) {
      sc.ws();
      if (sc.p < sc.end && *sc.p == '}') {
        ++sc.p;
        break;
      }
      std::string pat;
      if (sc.p < sc.end && *sc.p == '"') {
        ++sc.p;
        const char* s = sc.p;
        while (sc.p < sc.end && *sc.p != '"') ++sc.p;
        pat.assign(s, sc.p);
        sc.expect('"');
      } else {
        const char* s = sc.p;
        while (sc.p < sc.end && *sc.p != ',' && *sc.p != ' ' && *sc.p != '}' && *sc.p != '\n') ++sc.p;
        pat.assign(s, sc.p);
      }
      q.patterns.emplace_back(pat);
      sc.ws();
      if (sc.p < sc.end && *sc.p == ',') ++sc.p;
    }
    qindex[name] = (int)model->questions.size();  // later definitions win (BTreeMap::from_iter)
    model->questions.push_back(std::move(q));
    sc.ws();
  }
  while (sc.p < sc.end) {
    sc.ws();
    if (sc.p >= sc.end) break;
    if (!sc.starts("{*}")) parse_fail("expected {*} in tree section");
    sc.p += 3;
    sc.ws();
    sc.expect('[');
    std::string st;
    while (sc.p < sc.end && *sc.p != ']') st.push_back(*sc.p++);
    sc.expect(']');
    sc.ws();
    Tree tree;
    tree.state = (size_t)strtol(st.c_str(), nullptr, 10);
    std::vector<RawNode> raw;
    if (sc.p < sc.end && *sc.p == '{') {
      ++sc.p;
      for (;;) {
        sc.ws();
        if (sc.p < sc.end && *sc.p == '}') {
          ++sc.p;
          break;
        }
        RawNode nd;
        nd.id = strtol(sc.token().c_str(), nullptr, 10);
        sc.ws();
        nd.question = sc.token();
        sc.ws();
        nd.no = parse_child(sc.token());  // third column = "no" branch (tree.rs:96)
        sc.ws();
        nd.yes = parse_child(sc.token());
        raw.push_back(nd);
      }
      // convert (parser/model/mod.rs:71-136): internal nodes keep file order, leaves follow.
      std::map<long, size_t> node_lut;
      for (size_t i = 0; i < raw.size(); ++i) node_lut[raw[i].id] = i;
      std::vector<long> pdfs;
      for (const auto& nd : raw) {
        if (!nd.yes.is_node) pdfs.push_back(nd.yes.id);
        if (!nd.no.is_node) pdfs.push_back(nd.no.id);
      }
      std::sort(pdfs.begin(), pdfs.end());
      auto resolve = [&](const RawChild& c) -> size_t {
        if (c.is_node) {
          auto it = node_lut.find(c.id);
          if (it == node_lut.end()) parse_fail("tree refers to unknown node id");
          return it->second;
        }
        auto it = std::lower_bound(pdfs.begin(), pdfs.end(), c.id);
        return (size_t)(it - pdfs.begin()) + raw.size();
      };
      for (const auto& nd : raw) {
        TreeNode tn;
        auto q = qindex.find(nd.question);
        if (q == qindex.end()) parse_fail("tree refers to unknown question " + nd.question);
        tn.question = q->second;
        tn.yes = resolve(nd.yes);
        tn.no = resolve(nd.no);
        tree.nodes.push_back(tn);
      }
      for (long id : pdfs) {
        TreeNode tn;
        tn.leaf = true;
        tn.pdf_index = (size_t)id;
        tree.nodes.push_back(tn);
      }
    } else {
      RawChild c = parse_child(sc.token());
      if (c.is_node) parse_fail("malformed single-leaf tree");
      TreeNode tn;
      tn.leaf = true;
      tn.pdf_index = (size_t)c.id;
      tree.nodes.push_back(tn);
    }
    model->trees.push_back(std::move(tree));
  }
}

Model parse_model(const uint8_t* data, size_t dlen, std::pair<size_t, size_t> tree_range,
                  std::pair<size_t, size_t> pdf_range, size_t pdf_len) {
  Model m;
  if (tree_range.second >= dlen || pdf_range.second >= dlen || tree_range.first > tree_range.second ||
      pdf_range.first > pdf_range.second)
    parse_fail("position out of range");
  parse_tree_section(data + tree_range.first, tree_range.second - tree_range.first + 1, &m);

  const uint8_t* p = data + pdf_range.first;
  const uint8_t* end = data + pdf_range.second + 1;
  size_t ntree = m.trees.size();
  if ((size_t)(end - p) < ntree * 4) parse_fail("Not enough data (pdf counts)");
  std::vector<uint32_t> npdf(ntree);
  for (size_t i = 0; i < ntree; ++i) {
    npdf[i] = (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
    p += 4;
  }
  m.pdf.resize(ntree);
  size_t half = pdf_len / 2;
  for (size_t t = 0; t < ntree; ++t) {
    m.pdf[t].resize(npdf[t]);
    for (uint32_t k = 0; k < npdf[t]; ++k) {
      if ((size_t)(end - p) < pdf_len * 4) parse_fail("Not enough data (pdf)");
      std::vector<double> lin(pdf_len);
      for (size_t i = 0; i < pdf_len; ++i) {
        uint32_t u = (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
        float f;
        memcpy(&f, &u, 4);
        lin[i] = (double)f;  // LE f32 widened to f64 (parser/model/mod.rs:49)
        p += 4;
      }
      ModelParameter& mp = m.pdf[t][k];
      mp.parameters.resize(half);
      for (size_t i = 0; i < half; ++i) mp.parameters[i] = {lin[i], lin[i + half]};
      if (pdf_len > half * 2) {
        mp.has_msd = true;
        mp.msd = lin[half * 2];
      }
    }
  }
  if (p != end) parse_fail("trailing bytes after pdf block");
  return m;
}

Window parse_window(const uint8_t* data, size_t len) {
  std::string s((const char*)data, len);
  char* e = nullptr;
  const char* c = s.c_str();
  long n = strtol(c, &e, 10);
  if (e == c || n < 0) parse_fail("bad window row");
  Window w;
  for (long i = 0; i < n; ++i) {
    const char* b = e;
    double v = strtod(b, &e);
    if (e == b) parse_fail("bad window coefficient");
    w.coefficients.push_back(v);
  }
  while (*e == ' ' || *e == '\n') ++e;
  if (*e != 0) parse_fail("trailing characters in window row");
  return w;
}

}  // namespace

std::shared_ptr<Voice> parse_htsvoice(const uint8_t* bytes, size_t len) {
  std::string all((const char*)bytes, len);
  auto find_tag = [&](const char* tag, size_t from) {
    size_t p = all.find(tag, from);
    if (p == std::string::npos) parse_fail(std::string("section ") + tag + " not found");
    return p;
  };
  size_t g = find_tag("[GLOBAL]\n", 0);
  size_t s = find_tag("\n[STREAM]\n", g);
  size_t po = find_tag("\n[POSITION]\n", s);
  size_t d = find_tag("\n[DATA]\n", po);
  for (size_t i = 0; i < g; ++i)
    if (all[i] != '\n') parse_fail("garbage before [GLOBAL]");
  Header hg = parse_header(all.substr(g + 9, s - (g + 9)));
  Header hs = parse_header(all.substr(s + 10, po - (s + 10)));
  Header hp = parse_header(all.substr(po + 12, d - (po + 12)));
  const uint8_t* data = bytes + d + 8;
  size_t dlen = len - (d + 8);

  auto voice = std::make_shared<Voice>();
  GlobalModelMetadata& md = voice->metadata;
  md.hts_voice_version = one_string(need(hg, "HTS_VOICE_VERSION"));
  md.sampling_frequency = to_usize(one_string(need(hg, "SAMPLING_FREQUENCY")));
  md.frame_period = to_usize(one_string(need(hg, "FRAME_PERIOD")));
  md.num_states = to_usize(one_string(need(hg, "NUM_STATES")));
  md.num_streams = to_usize(one_string(need(hg, "NUM_STREAMS")));
  md.stream_type = need(hg, "STREAM_TYPE");
  md.fullcontext_format = one_string(need(hg, "FULLCONTEXT_FORMAT"));
  md.fullcontext_version = one_string(need(hg, "FULLCONTEXT_VERSION"));
  if (hg.plain.count("GV_OFF_CONTEXT")) md.gv_off_context_patterns = hg.plain["GV_OFF_CONTEXT"];
  for (const auto& p : md.gv_off_context_patterns) md.gv_off_context.patterns.emplace_back(p);

  voice->duration_model = parse_model(data, dlen, to_range(one_string(need(hp, "DURATION_TREE"))),
                                      to_range(one_string(need(hp, "DURATION_PDF"))),
                                      md.num_states * 2);

  auto idx = [&](const Header& h, const std::string& key, const std::string& stream,
                 bool required) -> const std::vector<std::string>* {
    auto it = h.indexed.find(key);
    if (it == h.indexed.end()) {
      if (required) parse_fail("missing field " + key);
      return nullptr;
    }
    auto jt = it->second.find(stream);
    if (jt == it->second.end()) {
      if (required) parse_fail("missing field " + key + "[" + stream + "]");
      return nullptr;
    }
    return &jt->second;
  };

  for (const std::string& key : md.stream_type) {
    StreamModels sm;
    sm.metadata.vector_length = to_usize(one_string(*idx(hs, "VECTOR_LENGTH", key, true)));
    sm.metadata.num_windows = to_usize(one_string(*idx(hs, "NUM_WINDOWS", key, true)));
    sm.metadata.is_msd = to_bool(*idx(hs, "IS_MSD", key, true));
    if (auto* v = idx(hs, "USE_GV", key, false)) sm.metadata.use_gv = to_bool(*v);
    if (auto* v = idx(hs, "OPTION", key, false)) sm.metadata.option = *v;

    size_t pdf_len = sm.metadata.vector_length * sm.metadata.num_windows * 2 +
                     (sm.metadata.is_msd ? 1 : 0);
    sm.stream_model = parse_model(data, dlen, to_range(one_string(*idx(hp, "STREAM_TREE", key, true))),
                                  to_range(one_string(*idx(hp, "STREAM_PDF", key, true))), pdf_len);
    if (sm.metadata.use_gv) {
      auto* gt = idx(hp, "GV_TREE", key, false);
      auto* gp = idx(hp, "GV_PDF", key, false);
      if (!gt || !gp)
        throw Error(kErrParse, "USE_GV is true, but positions for GV is not set");
      sm.has_gv = true;
      sm.gv_model = parse_model(data, dlen, to_range(one_string(*gt)), to_range(one_string(*gp)),
                                sm.metadata.vector_length * 2);
    }
    for (const std::string& r : *idx(hp, "STREAM_WIN", key, true)) {
      auto range = to_range(r);
      if (range.second >= dlen || range.first > range.second) parse_fail("window out of range");
      sm.windows.windows.push_back(parse_window(data + range.first, range.second - range.first + 1));
    }
    voice->stream_models.push_back(std::move(sm));
  }
  return voice;
}

}  // namespace jb
