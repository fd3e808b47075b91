#include "model.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <sstream>
#include <stdexcept>

namespace lgb {

namespace {
const double kZeroThresholdD = 1e-35f;
inline double RoundTinyToZero(double v) { return (v >= -kZeroThresholdD && v <= kZeroThresholdD) ? 0.0 : v; }
std::string G17(double v) { char b[64]; snprintf(b, sizeof(b), "%.17g", v); return b; }
std::string Gs(double v) { char b[64]; snprintf(b, sizeof(b), "%g", v); return b; }

template <typename T, typename F>
std::string JoinN(const std::vector<T>& v, int n, F f) {
  std::string s;
  for (int i = 0; i < n; ++i) {
    if (i) s.push_back(' ');
    s += f(v[i]);
  }
  return s;
}
}  // namespace

void HostTree::Shrinkage(double rate) {
  for (int i = 0; i < num_leaves - 1; ++i) {
    leaf_value[i] = RoundTinyToZero(leaf_value[i] * rate);
    internal_value[i] = RoundTinyToZero(internal_value[i] * rate);
  }
  leaf_value[num_leaves - 1] = RoundTinyToZero(leaf_value[num_leaves - 1] * rate);
  shrinkage *= rate;
}

void HostTree::AddBias(double val) {
  for (int i = 0; i < num_leaves - 1; ++i) {
    leaf_value[i] = RoundTinyToZero(leaf_value[i] + val);
    internal_value[i] = RoundTinyToZero(internal_value[i] + val);
  }
  leaf_value[num_leaves - 1] = RoundTinyToZero(leaf_value[num_leaves - 1] + val);
  shrinkage = 1.0;
}

void HostTree::AsConstantTree(double val) {
  num_leaves = 1;
  shrinkage = 1.0;
  leaf_value.assign(1, val);
}

bool HostTree::CategoricalGoesLeft(int node, double fval) const {
  if (std::isnan(fval) || !(fval < 2147483648.0)) return false;
  const int iv = static_cast<int>(fval);
  if (iv < 0) return false;
  const int k = static_cast<int>(threshold[node]);
  const int words = cat_boundaries[k + 1] - cat_boundaries[k];
  const int w = iv / 32;
  if (w >= words) return false;
  return (cat_threshold[cat_boundaries[k] + w] >> (iv % 32)) & 1u;
}

int HostTree::GetLeaf(const double* row) const {
  if (num_leaves <= 1) return 0;
  int node = 0;
  while (node >= 0) {
    double fval = row[split_feature[node]];
    if (decision_type[node] & 1) { node = CategoricalGoesLeft(node, fval) ? left_child[node] : right_child[node]; continue; }
    const int mt = (decision_type[node] >> 2) & 3;
    const bool default_left = (decision_type[node] & 2) != 0;
    if (std::isnan(fval) && mt != 2) fval = 0.0;
    if ((mt == 1 && fval >= -kZeroThresholdD && fval <= kZeroThresholdD) || (mt == 2 && std::isnan(fval)))
      node = default_left ? left_child[node] : right_child[node];
    else
      node = fval <= threshold[node] ? left_child[node] : right_child[node];
  }
  return ~node;
}

std::string HostTree::ToString() const {
  auto fi = [](int v) { return std::to_string(v); };
  auto fd = [](int8_t v) { return std::to_string(static_cast<int>(v)); };
  auto f17 = [](double v) { return G17(v); };
  auto fg = [](float v) { return Gs(v); };
  auto fu = [](uint32_t v) { return std::to_string(v); };
  std::string s = "num_leaves=" + std::to_string(num_leaves) + "\nnum_cat=" + std::to_string(num_leaves > 1 ? num_cat : 0) + "\n";
  if (num_leaves > 1) {
    const int ni = num_leaves - 1;
    s += "split_feature=" + JoinN(split_feature, ni, fi) + "\n";
    s += "split_gain=" + JoinN(split_gain, ni, fg) + "\n";
    s += "threshold=" + JoinN(threshold, ni, f17) + "\n";
    s += "decision_type=" + JoinN(decision_type, ni, fd) + "\n";
    s += "left_child=" + JoinN(left_child, ni, fi) + "\n";
    s += "right_child=" + JoinN(right_child, ni, fi) + "\n";
    s += "leaf_value=" + JoinN(leaf_value, num_leaves, f17) + "\n";
    s += "leaf_weight=" + JoinN(leaf_weight, num_leaves, f17) + "\n";
    s += "leaf_count=" + JoinN(leaf_count, num_leaves, fi) + "\n";
    s += "internal_value=" + JoinN(internal_value, ni, f17) + "\n";
    s += "internal_weight=" + JoinN(internal_weight, ni, f17) + "\n";
    s += "internal_count=" + JoinN(internal_count, ni, fi) + "\n";
    if (num_cat > 0) {
      s += "cat_boundaries=" + JoinN(cat_boundaries, num_cat + 1, fi) + "\n";
      s += "cat_threshold=" + JoinN(cat_threshold, static_cast<int>(cat_threshold.size()), fu) + "\n";
    }
  } else {
    s += "leaf_value=" + G17(leaf_value[0]) + "\n";
  }
  s += "is_linear=0\nshrinkage=" + Gs(shrinkage) + "\n\n";
  return s;
}

std::string ModelToString(const ModelHeader& h, const std::vector<HostTree>& trees, int start_iteration, int num_iteration) {
  const int tpi = std::max(1, h.num_tree_per_iteration);
  int first = std::min(std::max(0, start_iteration) * tpi, static_cast<int>(trees.size()));
  int last = static_cast<int>(trees.size());
  if (num_iteration > 0) last = std::min(last, first + num_iteration * tpi);
  std::string s = "tree\nversion=v3\nnum_class=" + std::to_string(h.num_class) + "\nnum_tree_per_iteration=" + std::to_string(tpi) + "\nlabel_index=0\n";
  s += "max_feature_idx=" + std::to_string(h.max_feature_idx) + "\n";
  if (!h.objective.empty()) s += "objective=" + h.objective + "\n";
  if (h.average_output) s += "average_output\n";
  s += "feature_names=";
  for (size_t j = 0; j < h.feature_names.size(); ++j) { if (j) s.push_back(' '); s += h.feature_names[j]; }
  s += "\nfeature_infos=";
  for (size_t j = 0; j < h.feature_infos.size(); ++j) { if (j) s.push_back(' '); s += h.feature_infos[j]; }
  s += "\n";
  std::vector<std::string> blocks;
  for (int i = first; i < last; ++i) blocks.push_back("Tree=" + std::to_string(i - first) + "\n" + trees[i].ToString() + "\n");
  s += "tree_sizes=";
  for (size_t i = 0; i < blocks.size(); ++i) { if (i) s.push_back(' '); s += std::to_string(blocks[i].size()); }
  s += "\n\n";
  for (auto& b : blocks) s += b;
  s += "end of trees\n\n";
  std::vector<int> imp(h.feature_names.size(), 0);
  for (int i = first; i < last; ++i)
    for (int k = 0; k < trees[i].num_leaves - 1; ++k) imp[trees[i].split_feature[k]]++;
  std::vector<int> order;
  for (size_t j = 0; j < imp.size(); ++j) if (imp[j] > 0) order.push_back(static_cast<int>(j));
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return imp[a] > imp[b]; });
  s += "feature_importances:\n";
  for (int j : order) s += h.feature_names[j] + "=" + std::to_string(imp[j]) + "\n";
  s += "\nparameters:\n";
  for (auto& kv : h.parameters) s += "[" + kv.first + ": " + kv.second + "]\n";
  s += "end of parameters\n\npandas_categorical:null\n";
  return s;
}

namespace {
std::vector<std::string> SplitWs(const std::string& v) {
  std::vector<std::string> out;
  std::istringstream is(v);
  std::string t;
  while (is >> t) out.push_back(t);
  return out;
}
template <typename T, typename F>
void ParseList(const std::map<std::string, std::string>& kv, const char* key, size_t n, std::vector<T>* out, F conv) {
  auto it = kv.find(key);
  if (it == kv.end()) throw std::runtime_error(std::string("Model string: tree block lacks '") + key + "'");
  const std::vector<std::string> tok = SplitWs(it->second);
  if (tok.size() != n) throw std::runtime_error(std::string("Model string: '") + key + "' has " + std::to_string(tok.size()) +
                                                " values, expected " + std::to_string(n));
  out->resize(n);
  for (size_t i = 0; i < n; ++i) (*out)[i] = conv(tok[i]);
}
}  // namespace

void ParseModelString(const std::string& text, ModelHeader* h, std::vector<HostTree>* trees) {
  *h = ModelHeader();
  trees->clear();
  std::istringstream is(text);
  std::string line;
  auto kv_of = [](const std::string& l, std::string* k, std::string* v) {
    const size_t eq = l.find('=');
    if (eq == std::string::npos) return false;
    *k = l.substr(0, eq); *v = l.substr(eq + 1);
    return true;
  };
  // ---- header
  bool saw_tree_word = false;
  std::map<std::string, std::string> cur;
  bool in_tree = false, in_params = false;
  auto finish_tree = [&]() {
    if (!in_tree) return;
    HostTree t;
    auto geti = [&](const char* k) { auto it = cur.find(k); if (it == cur.end()) throw std::runtime_error(std::string("Model string: tree block lacks '") + k + "'"); return atoi(it->second.c_str()); };
    t.num_leaves = geti("num_leaves");
    t.num_cat = cur.count("num_cat") ? atoi(cur["num_cat"].c_str()) : 0;
    if (t.num_cat < 0) throw std::runtime_error("Model string: num_cat < 0");
    if (cur.count("is_linear") && atoi(cur["is_linear"].c_str()) != 0) throw std::runtime_error("Model string: linear trees are not supported by this library");
    if (t.num_leaves < 1) throw std::runtime_error("Model string: num_leaves < 1");
    auto toi = [](const std::string& x) { return atoi(x.c_str()); };
    auto tod = [](const std::string& x) { return strtod(x.c_str(), nullptr); };
    auto tof = [](const std::string& x) { return strtof(x.c_str(), nullptr); };
    auto to8 = [](const std::string& x) { return static_cast<int8_t>(atoi(x.c_str())); };
    const size_t nl = static_cast<size_t>(t.num_leaves), ni = nl - 1;
    if (t.num_leaves > 1) {
      ParseList(cur, "split_feature", ni, &t.split_feature, toi);
      ParseList(cur, "split_gain", ni, &t.split_gain, tof);
      ParseList(cur, "threshold", ni, &t.threshold, tod);
      ParseList(cur, "decision_type", ni, &t.decision_type, to8);
      ParseList(cur, "left_child", ni, &t.left_child, toi);
      ParseList(cur, "right_child", ni, &t.right_child, toi);
      ParseList(cur, "leaf_value", nl, &t.leaf_value, tod);
      if (cur.count("leaf_weight")) ParseList(cur, "leaf_weight", nl, &t.leaf_weight, tod); else t.leaf_weight.assign(nl, 0.0);
      if (cur.count("leaf_count")) ParseList(cur, "leaf_count", nl, &t.leaf_count, toi); else t.leaf_count.assign(nl, 0);
      if (cur.count("internal_value")) ParseList(cur, "internal_value", ni, &t.internal_value, tod); else t.internal_value.assign(ni, 0.0);
      if (cur.count("internal_weight")) ParseList(cur, "internal_weight", ni, &t.internal_weight, tod); else t.internal_weight.assign(ni, 0.0);
      if (cur.count("internal_count")) ParseList(cur, "internal_count", ni, &t.internal_count, toi); else t.internal_count.assign(ni, 0);
      if (t.num_cat > 0) {
        auto tou = [](const std::string& x) { return static_cast<uint32_t>(strtoul(x.c_str(), nullptr, 10)); };
        ParseList(cur, "cat_boundaries", static_cast<size_t>(t.num_cat) + 1, &t.cat_boundaries, toi);
        if (t.cat_boundaries[0] != 0) throw std::runtime_error("Model string: cat_boundaries must start at 0");
        for (int k = 0; k < t.num_cat; ++k)
          if (t.cat_boundaries[k + 1] < t.cat_boundaries[k]) throw std::runtime_error("Model string: cat_boundaries must not decrease");
        ParseList(cur, "cat_threshold", static_cast<size_t>(t.cat_boundaries[t.num_cat]), &t.cat_threshold, tou);
      }
      for (size_t i = 0; i < ni; ++i) {
        if ((t.decision_type[i] & 1) && !(t.threshold[i] >= 0 && t.threshold[i] < t.num_cat))
          throw std::runtime_error("Model string: tree " + std::to_string(trees->size()) + " has a categorical node without a category set");
        const int l = t.left_child[i], r = t.right_child[i];
        auto ok = [&](int c) { return c >= 0 ? c < static_cast<int>(ni) : ~c < t.num_leaves; };
        if (!ok(l) || !ok(r) || t.split_feature[i] < 0 || t.split_feature[i] > h->max_feature_idx)
          throw std::runtime_error("Model string: tree " + std::to_string(trees->size()) + " has an out-of-range child or feature index");
      }
      t.split_feature_inner = t.split_feature;      // a loaded model has no bin mappers: inner indices and bin thresholds are unknown
      t.threshold_in_bin.assign(ni, 0);
    } else {
      ParseList(cur, "leaf_value", 1, &t.leaf_value, tod);
    }
    t.shrinkage = cur.count("shrinkage") ? strtod(cur["shrinkage"].c_str(), nullptr) : 1.0;
    trees->push_back(std::move(t));
    cur.clear();
    in_tree = false;
  };
  while (std::getline(is, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (line.empty()) continue;
    if (in_params) {
      if (line == "end of parameters") { in_params = false; continue; }
      if (line.size() > 2 && line.front() == '[' && line.back() == ']') {
        const size_t c = line.find(": ");
        if (c != std::string::npos) h->parameters.emplace_back(line.substr(1, c - 1), line.substr(c + 2, line.size() - c - 3));
      }
      continue;
    }
    if (line == "tree") { saw_tree_word = true; continue; }
    if (line == "average_output" && !in_tree) { h->average_output = true; continue; }
    if (line == "end of trees") { finish_tree(); continue; }
    if (line == "parameters:") { finish_tree(); in_params = true; continue; }
    if (line == "feature_importances:" || line.rfind("pandas_categorical", 0) == 0) { finish_tree(); continue; }
    std::string k, v;
    if (!kv_of(line, &k, &v)) continue;
    if (k == "Tree") { finish_tree(); in_tree = true; continue; }
    if (in_tree) { cur[k] = v; continue; }
    if (k == "max_feature_idx") h->max_feature_idx = atoi(v.c_str());
    else if (k == "objective") h->objective = v;
    else if (k == "feature_names") h->feature_names = SplitWs(v);
    else if (k == "feature_infos") h->feature_infos = SplitWs(v);
    else if (k == "num_class") { h->num_class = atoi(v.c_str()); if (h->num_class < 1) throw std::runtime_error("Model string: bad num_class"); }
    else if (k == "num_tree_per_iteration") { h->num_tree_per_iteration = atoi(v.c_str()); if (h->num_tree_per_iteration < 1) throw std::runtime_error("Model string: bad num_tree_per_iteration"); }
  }
  finish_tree();
  if (!saw_tree_word) throw std::runtime_error("Model string does not start with 'tree'");
}

namespace {
double AvoidInf(double x) { return std::isnan(x) ? 0.0 : x >= 1e300 ? 1e300 : x <= -1e300 ? -1e300 : x; }

void NodeToJson(const HostTree& t, int index, std::string* s) {
  if (index >= 0) {
    const int dt = t.decision_type[index];
    const int mt = (dt >> 2) & 3;
    *s += "{\"split_index\":" + std::to_string(index) + ",\n";
    *s += "\"split_feature\":" + std::to_string(t.split_feature[index]) + ",\n";
    *s += "\"split_gain\":" + G17(AvoidInf(t.split_gain[index])) + ",\n";
    if (dt & 1) {      // Tree::NodeToJSON: the categories joined by "||", decision "=="
      const int k = static_cast<int>(t.threshold[index]);
      std::string cats;
      for (int w = t.cat_boundaries[k]; w < t.cat_boundaries[k + 1]; ++w)
        for (int j = 0; j < 32; ++j)
          if ((t.cat_threshold[w] >> j) & 1u) { if (!cats.empty()) cats += "||"; cats += std::to_string((w - t.cat_boundaries[k]) * 32 + j); }
      *s += "\"threshold\":\"" + cats + "\",\n";
      *s += "\"decision_type\":\"==\",\n";
    } else {
      *s += "\"threshold\":" + G17(AvoidInf(t.threshold[index])) + ",\n";
      *s += "\"decision_type\":\"<=\",\n";
    }
    *s += std::string("\"default_left\":") + ((dt & 2) ? "true" : "false") + ",\n";
    *s += std::string("\"missing_type\":\"") + (mt == 0 ? "None" : mt == 1 ? "Zero" : "NaN") + "\",\n";
    *s += "\"internal_value\":" + G17(t.internal_value[index]) + ",\n";
    *s += "\"internal_weight\":" + G17(t.internal_weight[index]) + ",\n";
    *s += "\"internal_count\":" + std::to_string(t.internal_count[index]) + ",\n";
    *s += "\"left_child\":";
    NodeToJson(t, t.left_child[index], s);
    *s += ",\n\"right_child\":";
    NodeToJson(t, t.right_child[index], s);
    *s += "\n}";
  } else {
    const int leaf = ~index;
    *s += "{\"leaf_index\":" + std::to_string(leaf) + ",\n";
    *s += "\"leaf_value\":" + G17(t.leaf_value[leaf]) + ",\n";
    *s += "\"leaf_weight\":" + G17(t.leaf_weight[leaf]) + ",\n";
    *s += "\"leaf_count\":" + std::to_string(t.leaf_count[leaf]) + "\n}";
  }
}

std::string JsonEscape(const std::string& v) {
  std::string o;
  for (char c : v) { if (c == '"' || c == '\\') o.push_back('\\'); o.push_back(c); }
  return o;
}
}  // namespace

std::string ModelToJson(const ModelHeader& h, const std::vector<HostTree>& trees, int start_iteration, int num_iteration,
                        const std::vector<double>& importance) {
  const int tpi = std::max(1, h.num_tree_per_iteration);
  int first = std::min(std::max(0, start_iteration) * tpi, static_cast<int>(trees.size()));
  int last = static_cast<int>(trees.size());
  if (num_iteration > 0) last = std::min(last, first + num_iteration * tpi);
  std::string s = "{\"name\":\"tree\",\n\"version\":\"v3\",\n\"num_class\":" + std::to_string(h.num_class) + ",\n\"num_tree_per_iteration\":" +
                  std::to_string(tpi) + ",\n\"label_index\":0,\n";
  s += "\"max_feature_idx\":" + std::to_string(h.max_feature_idx) + ",\n";
  if (!h.objective.empty()) s += "\"objective\":\"" + JsonEscape(h.objective) + "\",\n";
  s += std::string("\"average_output\":") + (h.average_output ? "true" : "false") + ",\n\"feature_names\":[";
  for (size_t j = 0; j < h.feature_names.size(); ++j) { if (j) s += ","; s += "\"" + JsonEscape(h.feature_names[j]) + "\""; }
  s += "],\n\"monotone_constraints\":[],\n\"feature_infos\":{";
  bool first_obj = true;
  for (size_t j = 0; j < h.feature_infos.size() && j < h.feature_names.size(); ++j) {
    const std::string& fi = h.feature_infos[j];
    const size_t colon = fi.find(':');
    if (fi == "none" || colon == std::string::npos) continue;      // a trivial feature
    if (fi[0] != '[') {      // categorical: the categories of its bins, "-1:3:1:..."
      std::vector<long long> vals;
      std::stringstream ss(fi);
      std::string tok;
      while (std::getline(ss, tok, ':')) if (!tok.empty()) vals.push_back(atoll(tok.c_str()));
      if (vals.empty()) continue;
      if (!first_obj) s += ",";
      first_obj = false;
      s += "\"" + JsonEscape(h.feature_names[j]) + "\":{\"min_value\":" + std::to_string(*std::min_element(vals.begin(), vals.end())) +
           ",\"max_value\":" + std::to_string(*std::max_element(vals.begin(), vals.end())) + ",\"values\":[";
      for (size_t k = 0; k < vals.size(); ++k) { if (k) s += ","; s += std::to_string(vals[k]); }
      s += "]}";
      continue;
    }
    if (fi.size() < 5) continue;
    if (!first_obj) s += ",";
    first_obj = false;
    s += "\"" + JsonEscape(h.feature_names[j]) + "\":{\"min_value\":" + fi.substr(1, colon - 1) + ",\"max_value\":" +
         fi.substr(colon + 1, fi.size() - colon - 2) + ",\"values\":[]}";
  }
  s += "},\n\"tree_info\":[";
  for (int i = first; i < last; ++i) {
    const HostTree& t = trees[i];
    if (i > first) s += ",";
    s += "{\"tree_index\":" + std::to_string(i) + ",\"num_leaves\":" + std::to_string(t.num_leaves) + ",\n\"num_cat\":" + std::to_string(t.num_leaves > 1 ? t.num_cat : 0) + ",\n";
    s += "\"shrinkage\":" + G17(t.shrinkage) + ",\n\"tree_structure\":";
    if (t.num_leaves <= 1) s += "{\"leaf_value\":" + G17(t.leaf_value.empty() ? 0.0 : t.leaf_value[0]) + "}";
    else NodeToJson(t, 0, &s);
    s += "}";
  }
  s += "],\n\"feature_importances\":{";
  std::vector<int> order;
  for (size_t j = 0; j < importance.size() && j < h.feature_names.size(); ++j) if (importance[j] > 0) order.push_back(static_cast<int>(j));
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return importance[a] > importance[b]; });
  for (size_t k = 0; k < order.size(); ++k) {
    if (k) s += ",";
    s += "\"" + JsonEscape(h.feature_names[order[k]]) + "\":" + G17(importance[order[k]]);
  }
  s += "}\n}\n";
  return s;
}

std::string FormatG17(double v) { return G17(v); }
std::string FormatG(double v) { return Gs(v); }

}  // namespace lgb
