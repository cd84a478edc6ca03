// graph.h -- in-memory form of the "DUSPGRAPH 1" circuit description
// (format: dusp_b200/graphblob.py).  Host-only.
#pragma once
#include <map>
#include <string>
#include <vector>

namespace dusp {

struct GPort {
  std::string name;
  bool mono = false;      // Piglet.js:13-17 exposeAsMono
  bool nodata = false;    // Unit.js:88-98 chain ports
  bool connected = false;
  int src_unit = -1, src_outlet = -1;
  std::vector<double> consts;  // unconnected inlet: per-channel chunk constant (Inlet.js:76-93)
  int nch = 1;                 // outlets: initial channel count
};

struct GUnit {
  std::string kind, label;
  std::map<std::string, double> P;
  std::map<std::string, std::string> S;
  std::vector<GPort> in, out;
  double num(const std::string &k, double d) const { auto it = P.find(k); return it == P.end() ? d : it->second; }
  bool has(const std::string &k) const { return P.count(k) != 0; }
  std::string str(const std::string &k, const std::string &d) const { auto it = S.find(k); return it == S.end() ? d : it->second; }
};

struct GCircleBuffer { int nch; long long len; };
// one assignment of `target.trigger()` for a (Sporadic)Retriggerer unit: TG <unit> <target> <prop> <value>
struct GTrigger { int unit, target; std::string prop; double value; };

struct Graph {
  std::vector<GUnit> units;
  std::vector<GCircleBuffer> cbs;
  std::vector<int> order;
  std::vector<GTrigger> triggers;
  bool keep_all = false;  // KEEP record: a graph edit follows, keep every outlet's last chunk (compile.cc)
  int warmup_tick = -1;   // WARMUP k record: compile the circuit as its k-th tick sees it (channel counts still growing)
  int root_unit = -1, root_outlet = 0;
};

// Throws std::runtime_error on malformed input.
Graph parse_graph(const char *blob, size_t len);

}  // namespace dusp
