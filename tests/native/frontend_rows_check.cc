// Host front-end, index mode: the StreamTable rows build_utterance sends to the device (one Model::get_all lookup per
// label and voice, the label hashed once) against Model::get_index state by state (voice/model.rs:50-67) on a second,
// separately parsed copy of the voice - its caches are its own, so its first pass really walks the trees. Two passes:
// cold caches, then warm. Usage: frontend_rows_check voice.htsvoice utterances.txt (blank line between utterances).
#include <cstdio>
#include <fstream>
#include <iterator>

#include "frontend.h"

using namespace jb;

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::ifstream vf(argv[1], std::ios::binary);
  std::vector<uint8_t> bytes((std::istreambuf_iterator<char>(vf)), std::istreambuf_iterator<char>());
  VoiceSet vs({parse_htsvoice(bytes.data(), bytes.size())});
  auto other = parse_htsvoice(bytes.data(), bytes.size());
  Condition cond;
  cond.load_model(vs);
  const std::vector<StreamTable> tables = build_stream_tables(vs);
  std::ifstream lf(argv[2]);
  std::vector<std::vector<std::string>> utts(1);
  for (std::string line; std::getline(lf, line);) {
    if (!line.empty()) utts.back().push_back(line);
    else if (!utts.back().empty()) utts.emplace_back();
  }
  if (utts.back().empty()) utts.pop_back();
  const size_t ns = vs.global_metadata().num_states;
  size_t differ = 0, n = 0, frames = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (const auto& u : utts) {
      const UtterancePack p = build_utterance(vs, cond, u, &tables);
      frames += p.nframes;
      if (p.nframes != utterance_frames(vs, cond, u)) ++differ;
      for (size_t s = 0; s < p.streams.size(); ++s) {
        if (!tables[s].regular) continue;
        const Model& m = other->stream_models[s].stream_model;
        if (p.streams[s].rows.size() != p.nstates) ++differ;
        for (size_t i = 0; i < p.nstates && i < p.streams[s].rows.size(); ++i, ++n) {
          long ti = -1, pi = -1;
          m.get_index(2 + i % ns, u[i / ns], &ti, &pi);
          if (ti < 2 || pi < 1) {
            ++differ;
            continue;
          }
          const size_t want = tables[s].voice_base[0] + tables[s].tree_base[0][(size_t)ti - 2] + (size_t)pi - 1;
          differ += (int32_t)want != p.streams[s].rows[i];
        }
      }
    }
  // a label no tree has seen: every question answers "no" and the walk still ends in a leaf, twice the same
  {
    const std::string odd = "x^x-zz+x=x/A:0+0+0/B:xx-xx_xx/C:xx_xx+xx/D:xx+xx_xx/E:xx_xx!xx_xx-xx/F:xx_xx#xx_xx@xx_xx|xx_xx/G:xx_xx%xx_xx_xx/H:xx_xx/I:xx-xx@xx+xx&xx-xx|xx+xx/J:xx_xx/K:0+0-0";
    const Model& a = vs.voices[0]->stream_models[0].stream_model;
    const Model& b = other->stream_models[0].stream_model;
    std::vector<int32_t> all(a.trees.size(), 0), again(a.trees.size(), 0);
    a.get_all(odd, label_hash(odd), all.data());
    a.get_all(odd, label_hash(odd), again.data());
    for (size_t t = 0; t < a.trees.size(); ++t, ++n) {
      long ti, pi;
      b.get_index(a.trees[t].state, odd, &ti, &pi);
      differ += all[t] != (int32_t)pi || again[t] != all[t] || ti != (long)t + 2;
    }
  }
  printf("%zu rows compared, %zu differ, %zu frames\n", n, differ, frames);
  return differ != 0;
}
