// graph.cc -- parser for the "DUSPGRAPH 1" circuit description.
#include "graph.h"

#include <cstdlib>
#include <sstream>
#include <stdexcept>

namespace dusp {

static std::vector<std::string> split(const std::string &line) {
  std::vector<std::string> t;
  std::istringstream ss(line);
  std::string w;
  while (ss >> w) t.push_back(w);
  return t;
}

Graph parse_graph(const char *blob, size_t len) {
  Graph g;
  std::string text(blob, len);
  std::istringstream in(text);
  std::string line;
  bool header = false, has_order = false;
  GUnit *u = nullptr;
  auto need = [&](const std::vector<std::string> &t, size_t n) {
    if (t.size() < n) throw std::runtime_error("malformed graph blob line: " + line);
  };
  while (std::getline(in, line)) {
    auto t = split(line);
    if (t.empty() || t[0][0] == '#') continue;
    const std::string &tag = t[0];
    if (tag == "DUSPGRAPH") { need(t, 2); if (t[1] != "1") throw std::runtime_error("unsupported DUSPGRAPH version"); header = true; continue; }
    if (!header) throw std::runtime_error("graph blob does not start with DUSPGRAPH 1");
    if (tag == "END") break;
    if (tag == "CB") { need(t, 3); g.cbs.push_back({atoi(t[1].c_str()), atoll(t[2].c_str())}); continue; }
    if (tag == "U") { need(t, 2); g.units.emplace_back(); u = &g.units.back(); u->kind = t[1]; u->label = t.size() > 2 ? t[2] : ""; continue; }
    if (tag == "EV") continue;  // events after the first chunk: applied by the host as pokes (dusp_session_poke)
    if (tag == "KEEP") { g.keep_all = true; continue; }
    if (tag == "WARMUP") { need(t, 2); g.warmup_tick = atoi(t[1].c_str()); continue; }
    if (tag == "TG") { need(t, 5); g.triggers.push_back({atoi(t[1].c_str()), atoi(t[2].c_str()), t[3], strtod(t[4].c_str(), nullptr)}); continue; }
    if (tag == "ORDER") { for (size_t i = 1; i < t.size(); i++) g.order.push_back(atoi(t[i].c_str())); has_order = true; continue; }
    if (tag == "ROOT") { need(t, 3); g.root_unit = atoi(t[1].c_str()); g.root_outlet = atoi(t[2].c_str()); continue; }
    if (!u) throw std::runtime_error("graph blob: property before any unit");
    if (tag == "P") { need(t, 3); u->P[t[1]] = strtod(t[2].c_str(), nullptr); }
    else if (tag == "S") { need(t, 3); u->S[t[1]] = t[2]; }
    else if (tag == "I") {
      need(t, 4);
      GPort p; p.name = t[1]; p.mono = atoi(t[2].c_str()) != 0;
      if (t[3] == "C") {
        need(t, 5);
        int n = atoi(t[4].c_str());
        need(t, 5 + (size_t)n);
        for (int i = 0; i < n; i++) p.consts.push_back(strtod(t[5 + i].c_str(), nullptr));
        if (p.consts.empty()) p.consts.push_back(0.0);
      } else if (t[3] == "K" || t[3] == "N") {
        need(t, 6);
        p.connected = true; p.nodata = (t[3] == "N");
        p.src_unit = atoi(t[4].c_str()); p.src_outlet = atoi(t[5].c_str());
      } else throw std::runtime_error("graph blob: bad inlet mode " + t[3]);
      u->in.push_back(p);
    } else if (tag == "O") {
      need(t, 4);
      GPort p; p.name = t[1]; p.mono = atoi(t[2].c_str()) != 0; p.nch = atoi(t[3].c_str());
      p.nodata = p.name.rfind("chain", 0) == 0;
      u->out.push_back(p);
    } else throw std::runtime_error("graph blob: unknown record " + tag);
  }
  if (g.units.empty()) throw std::runtime_error("renderAudioBuffer expects an outlet");  // renderChannelData.js:13
  if (!has_order) throw std::runtime_error("graph blob has no ORDER record (circuit.units)");
  if (g.root_unit < 0 || g.root_unit >= (int)g.units.size() || g.root_outlet < 0 ||
      g.root_outlet >= (int)g.units[g.root_unit].out.size())
    throw std::runtime_error("renderAudioBuffer expects an outlet");  // renderChannelData.js:17
  for (int i : g.order)
    if (i < 0 || i >= (int)g.units.size()) throw std::runtime_error("graph blob: ORDER names a missing unit");
  for (auto &tg : g.triggers)
    if (tg.unit < 0 || tg.unit >= (int)g.units.size() || tg.target < 0 || tg.target >= (int)g.units.size())
      throw std::runtime_error("graph blob: TG names a missing unit");
  for (auto &un : g.units)
    for (auto &p : un.in)
      if (p.connected && (p.src_unit < 0 || p.src_unit >= (int)g.units.size() || p.src_outlet < 0 ||
                          p.src_outlet >= (int)g.units[p.src_unit].out.size()))
        throw std::runtime_error("graph blob: inlet connected to a missing outlet");
  return g;
}

}  // namespace dusp
