// host/pm_compat.cpp — implementation of the libpointmatcher API stand-ins (see pm_compat.h) and of the host-side modules
// that stay on the CPU in the reference design too: RigidTransformation (for standalone use) and the TransformationCheckers
// (a12 in SURVEY 8a: "host, unchanged semantics").
#include "pm_compat.h"

#include <algorithm>
#include <cstring>
#include <mutex>

#include "cuda_modules.h"

namespace PointMatcherSupport {

// Parametrizable ctor (Parametrizable.cpp:170-192): defaults, bounds -> InvalidParameter with the reference's messages
Parametrizable::Parametrizable(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params)
    : className(className), parametersDoc(paramsDoc) {
  for (auto it = parametersDoc.begin(); it != parametersDoc.end(); ++it) {
    const std::string& paramName(it->name);
    auto paramIt = params.find(paramName);
    if (paramIt != params.end()) {
      const std::string& val(paramIt->second);
      if (it->comp) {
        if (it->comp(val, it->minValue))
          throw InvalidParameter("Value " + val + " of parameter " + paramName + " in class " + className + " is smaller than minimum admissible value " + it->minValue);
        if (it->comp(it->maxValue, val))
          throw InvalidParameter("Value " + val + " of parameter " + paramName + " in class " + className + " is larger than maximum admissible value " + it->maxValue);
      }
      parameters[paramName] = val;
    } else {
      parameters[paramName] = it->defaultValue;
    }
  }
}

std::string Parametrizable::getParamValueString(const std::string& paramName) {   // Parametrizable.cpp:199-207
  auto paramIt = parameters.find(paramName);
  if (paramIt == parameters.end()) throw InvalidParameter("Parameter " + paramName + " does not exist in class " + className);
  parametersUsed.insert(paramIt->first);
  return paramIt->second;
}

// ---------------------------------------------------------------------------------------------------------------
// YAML subset: block maps / block lists by indentation, "key: value", "- item", "key: {}" / "[]", comments, bare scalars.
// Enough for every ICP chain file of the reference (examples/data/*.yaml, icp_data/*.yaml, param/icp.yaml).
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct Line { int indent; std::string text; };

std::string stripComment(const std::string& s) {
  bool inS = false, inD = false;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '\'' && !inD) inS = !inS;
    else if (s[i] == '"' && !inS) inD = !inD;
    else if (s[i] == '#' && !inS && !inD && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
  }
  return s;
}
std::string trim(const std::string& s) {
  size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
  return a == std::string::npos ? "" : s.substr(a, b - a + 1);
}
std::string unquote(const std::string& s) {
  if (s.size() >= 2 && ((s.front() == '"' && s.back() == '"') || (s.front() == '\'' && s.back() == '\''))) return s.substr(1, s.size() - 2);
  return s;
}
YamlNode scalarNode(const std::string& v) {
  YamlNode n;
  const std::string t = trim(v);
  if (t.empty() || t == "~" || t == "null") return n;
  if (t == "{}") { n.kind = YamlNode::Map; return n; }
  if (t == "[]") { n.kind = YamlNode::List; return n; }
  n.kind = YamlNode::Scalar;
  n.scalar = unquote(t);
  return n;
}
YamlNode parseBlock(const std::vector<Line>& L, size_t& i, int indent);

// value after "key:" — inline scalar or a nested block on the following, deeper-indented lines
YamlNode parseValue(const std::vector<Line>& L, size_t& i, int parentIndent, const std::string& inlineText) {
  if (!trim(inlineText).empty()) return scalarNode(inlineText);
  if (i < L.size() && L[i].indent > parentIndent) return parseBlock(L, i, L[i].indent);
  // YAML allows a list at the SAME indent as its parent key ("key:\n- a")
  if (i < L.size() && L[i].indent == parentIndent && L[i].text.compare(0, 1, "-") == 0) return parseBlock(L, i, parentIndent);
  return YamlNode();
}

YamlNode parseBlock(const std::vector<Line>& L, size_t& i, int indent) {
  YamlNode node;
  while (i < L.size() && L[i].indent == indent) {
    const std::string& t = L[i].text;
    if (t[0] == '-' && (t.size() == 1 || t[1] == ' ')) {
      if (node.kind == YamlNode::Map) break;   // a list at the parent's indent ends this map
      node.kind = YamlNode::List;
      const std::string rest = trim(t.substr(1));
      const size_t firstContent = t.find_first_not_of(' ', 1);
      const int itemIndent = indent + (int)(firstContent == std::string::npos ? 2 : firstContent);   // column of the item's content
      ++i;
      if (rest.empty()) {
        node.list.push_back((i < L.size() && L[i].indent > indent) ? parseBlock(L, i, L[i].indent) : YamlNode());
      } else {
        const size_t colon = rest.find(':');
        const bool isKey = colon != std::string::npos && (colon + 1 == rest.size() || rest[colon + 1] == ' ');
        if (!isKey) {
          node.list.push_back(scalarNode(rest));
        } else {   // "- Name:" [+ params on deeper lines]  or "- Name: value"
          YamlNode item;
          item.kind = YamlNode::Map;
          const std::string key = unquote(trim(rest.substr(0, colon)));
          YamlNode val = parseValue(L, i, itemIndent, rest.substr(colon + 1));
          item.map.push_back({key, val});
          // further keys of the same list item sit at itemIndent
          while (i < L.size() && L[i].indent == itemIndent && !(L[i].text[0] == '-' && (L[i].text.size() == 1 || L[i].text[1] == ' '))) {
            const std::string& u = L[i].text;
            const size_t c2 = u.find(':');
            if (c2 == std::string::npos) throw ConfigurationError("YAML: expected 'key: value' in '" + u + "'");
            const std::string k2 = unquote(trim(u.substr(0, c2)));
            ++i;
            item.map.push_back({k2, parseValue(L, i, itemIndent, u.substr(c2 + 1))});
          }
          node.list.push_back(item);
        }
      }
    } else {
      if (node.kind == YamlNode::List) break;
      const size_t colon = t.find(':');
      if (colon == std::string::npos || !(colon + 1 == t.size() || t[colon + 1] == ' ')) {
        if (node.kind == YamlNode::Null && i + 1 >= L.size()) { ++i; return scalarNode(t); }   // bare scalar block
        if (node.kind == YamlNode::Null) { ++i; return scalarNode(t); }
        throw ConfigurationError("YAML: expected 'key: value' in '" + t + "'");
      }
      node.kind = YamlNode::Map;
      const std::string key = unquote(trim(t.substr(0, colon)));
      ++i;
      node.map.push_back({key, parseValue(L, i, indent, t.substr(colon + 1))});
    }
  }
  return node;
}
}  // namespace

YamlNode parseYaml(const std::string& text) {
  std::vector<Line> lines;
  std::istringstream is(text);
  std::string raw;
  while (std::getline(is, raw)) {
    std::string s = stripComment(raw);
    if (trim(s).empty() || trim(s) == "---") continue;
    int indent = 0;
    while (indent < (int)s.size() && s[indent] == ' ') ++indent;
    if (indent < (int)s.size() && s[indent] == '\t') throw ConfigurationError("YAML: tabs are not allowed for indentation");
    lines.push_back({indent, trim(s)});
  }
  size_t i = 0;
  if (lines.empty()) return YamlNode();
  YamlNode root = parseBlock(lines, i, lines[0].indent);
  if (i != lines.size()) throw ConfigurationError("YAML: inconsistent indentation near '" + lines[i].text + "'");
  return root;
}

// getNameParamsFromYAML (Registrar.cpp:7-26): a one-key map is name + string params, anything else is a bare name
void getNameParamsFromYAML(const YamlNode& module, std::string& name, Parametrizable::Parameters& params) {
  if (module.kind != YamlNode::Map || module.size() != 1) {
    if (module.kind != YamlNode::Scalar) throw ConfigurationError("YAML: a module entry must be a name or a single-key map");
    name = module.scalar;
  } else {
    name = module.map[0].first;
    const YamlNode& p = module.map[0].second;
    for (auto& kv : p.map) {
      if (kv.second.kind != YamlNode::Scalar) throw ConfigurationError("YAML: parameter " + kv.first + " of " + name + " is not a scalar");
      params[kv.first] = kv.second.scalar;
    }
  }
}

}  // namespace PointMatcherSupport

using namespace PointMatcherSupport;

// ---------------------------------------------------------------------------------------------------------------
// DataPoints helpers
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
int PointMatcher<T>::DataPoints::getDescriptorStartingRow(const std::string& name, size_t* span) const {
  int row = 0;
  for (auto& l : descriptorLabels) {
    if (l.text == name) { if (span) *span = l.span; return row; }
    row += (int)l.span;
  }
  throw InvalidField("Field " + name + " not found");
}

template <typename T>
int PointMatcher<T>::DataPoints::allocateDescriptor(const std::string& name, size_t span) {   // DataPoints.cpp:988-1027
  int row = 0;
  for (auto& l : descriptorLabels) {
    if (l.text == name) {
      if (l.span != span) throw InvalidField("The existing field " + name + " has dimension different than requested");
      return row;
    }
    row += (int)l.span;
  }
  const int64_t n = features.cols(), oldRows = descriptors.rows();
  Matrix nd(oldRows + (int64_t)span, n, T(0));
  for (int64_t j = 0; j < (descriptors.cols() == n ? n : 0); ++j)
    for (int64_t i = 0; i < oldRows; ++i) nd(i, j) = descriptors(i, j);
  descriptors = nd;
  descriptorLabels.push_back(Label(name, span));
  return (int)oldRows;
}

template <typename T>
void PointMatcher<T>::DataPoints::conservativeResize(int64_t n) {
  features.conservativeResizeCols(n);
  if (descriptors.rows() != 0) descriptors.conservativeResizeCols(n);
  if (times.rows() != 0) times.conservativeResizeCols(n);
}

template <typename T>
T PointMatcher<T>::Matches::getDistsQuantile(const T quantile) const {   // Matches.cpp:61-87
  std::vector<T> values;
  values.reserve(dists.d.size());
  for (const T v : dists.d) if (v != std::numeric_limits<T>::infinity()) values.push_back(v);
  if (values.empty()) throw ConvergenceError("No matches available for computing distance quantiles");
  if (quantile < 0.0 || quantile > 1.0) throw ConvergenceError("Distance quantile of matches must lie in the range [0,1]");
  if (quantile == 1.0) return *std::max_element(values.begin(), values.end());
  std::nth_element(values.begin(), values.begin() + (size_t)(values.size() * quantile), values.end());
  return values[(size_t)(values.size() * quantile)];
}

// ---------------------------------------------------------------------------------------------------------------
// host-side modules
// ---------------------------------------------------------------------------------------------------------------
namespace {

typedef PM::Matrix Matrix;
typedef PM::TransformationParameters TP;

// 4x4 fp32 product, depth accumulated k = 0..3 (same contract as the device kernels and the C ABI)
TP mul4(const TP& a, const TP& b) {
  TP r(4, 4);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float acc = a(i, 0) * b(0, j);
      acc = acc + a(i, 1) * b(1, j);
      acc = acc + a(i, 2) * b(2, j);
      acc = acc + a(i, 3) * b(3, j);
      r(i, j) = acc;
    }
  return r;
}
void toRowMajor(const TP& T, float out[16]) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[4 * i + j] = T(i, j); }
TP fromRowMajor(const float in[16]) { TP T(4, 4); for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) T(i, j) = in[4 * i + j]; return T; }

// RigidTransformation (TransformationsImpl.cpp:61-114), host version for standalone use of the shim
struct RigidTransformation : public PM::Transformation {
  static const std::string description() { return "Rigid transformation."; }
  bool checkParameters(const TP& p) const override {
    const float det = p(0, 0) * (p(1, 1) * p(2, 2) - p(1, 2) * p(2, 1)) - p(0, 1) * (p(1, 0) * p(2, 2) - p(1, 2) * p(2, 0)) +
                      p(0, 2) * (p(1, 0) * p(2, 1) - p(1, 1) * p(2, 0));
    return !(std::fabs(1.f - det) > 0.001f);
  }
  void inPlaceCompute(const TP& p, PM::DataPoints& cloud) const override {
    if (!checkParameters(p)) throw PM::TransformationError("RigidTransformation: Error, rotation matrix is not orthogonal.");
    const int64_t n = cloud.features.cols();
    for (int64_t c = 0; c < n; ++c) {
      const float x = cloud.features(0, c), y = cloud.features(1, c), z = cloud.features(2, c), w = cloud.features(3, c);
      for (int r = 0; r < 4; ++r) cloud.features(r, c) = ((p(r, 0) * x + p(r, 1) * y) + p(r, 2) * z) + p(r, 3) * w;
    }
    int row = 0;
    for (auto& l : cloud.descriptorLabels) {
      if (l.text == "normals" || l.text == "observationDirections")
        for (int64_t c = 0; c < n; ++c) {
          const float a = cloud.descriptors(row, c), b = cloud.descriptors(row + 1, c), d = cloud.descriptors(row + 2, c);
          for (int r = 0; r < 3; ++r) cloud.descriptors(row + r, c) = (p(r, 0) * a + p(r, 1) * b) + p(r, 2) * d;
        }
      row += (int)l.span;
    }
  }
  PM::DataPoints compute(const PM::DataPoints& input, const TP& p) const override { PM::DataPoints o(input); inPlaceCompute(p, o); return o; }
};

void quatOf(const TP& T, float q[4]) {   // Eigen::Quaternion(Matrix3)
  float t = T(0, 0) + T(1, 1) + T(2, 2);
  if (t > 0.f) {
    t = std::sqrt(t + 1.f); q[0] = 0.5f * t; t = 0.5f / t;
    q[1] = (T(2, 1) - T(1, 2)) * t; q[2] = (T(0, 2) - T(2, 0)) * t; q[3] = (T(1, 0) - T(0, 1)) * t;
  } else {
    int i = 0;
    if (T(1, 1) > T(0, 0)) i = 1;
    if (T(2, 2) > T(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(T(i, i) - T(j, j) - T(k, k) + 1.f);
    q[1 + i] = 0.5f * t; t = 0.5f / t;
    q[0] = (T(k, j) - T(j, k)) * t; q[1 + j] = (T(j, i) + T(i, j)) * t; q[1 + k] = (T(k, i) + T(i, k)) * t;
  }
}
float quatDist(const float a[4], const float b[4]) {
  const float bw = b[0], bx = -b[1], by = -b[2], bz = -b[3];
  const float w = a[0] * bw - a[1] * bx - a[2] * by - a[3] * bz, x = a[0] * bx + a[1] * bw + a[2] * bz - a[3] * by;
  const float y = a[0] * by + a[2] * bw + a[3] * bx - a[1] * bz, z = a[0] * bz + a[3] * bw + a[1] * by - a[2] * bx;
  return 2.f * std::atan2(std::sqrt(x * x + y * y + z * z), std::fabs(w));
}

}  // namespace

namespace b200 {
// TransformationCheckersImpl.{h,cpp}: Counter (:58-76), Differential (:80-158), Bound (:162-225)
struct CounterTransformationChecker : public PM::TransformationChecker {
  struct MaxNumIterationsReached {};
  static const std::string description() { return "This checker stops the ICP loop after a certain number of iterations."; }
  static const PM::ParametersDoc availableParameters() {
    return {{"maxIterationCount", "maximum number of iterations ", "40", "0", "2147483647", &PM::Parametrizable::Comp<unsigned>}};
  }
  const unsigned maxIterationCount;
  unsigned count = 0;
  explicit CounterTransformationChecker(const PM::Parameters& params = PM::Parameters())
      : PM::TransformationChecker("CounterTransformationChecker", availableParameters(), params), maxIterationCount(get<unsigned>("maxIterationCount")) {}
  void init(const TP&, bool&) override { count = 0; }
  void check(const TP&, bool& iterate) override {
    ++count;
    if (count >= maxIterationCount) { iterate = false; throw MaxNumIterationsReached(); }
  }
};
struct DifferentialTransformationChecker : public PM::TransformationChecker {
  static const std::string description() { return "This checker stops the ICP loop when the relative motions of rotation and translation components are below a fix thresholds."; }
  static const PM::ParametersDoc availableParameters() {
    return {{"minDiffRotErr", "threshold for rotation error (radian)", "0.001", "0.", "6.2831854", &PM::Parametrizable::Comp<float>},
            {"minDiffTransErr", "threshold for translation error", "0.001", "0.", "inf", &PM::Parametrizable::Comp<float>},
            {"smoothLength", "number of iterations over which to average the differencial error", "3", "0", "2147483647", &PM::Parametrizable::Comp<unsigned>}};
  }
  const float minDiffRotErr, minDiffTransErr;
  const unsigned smoothLength;
  std::vector<float> rot, tr;
  explicit DifferentialTransformationChecker(const PM::Parameters& params = PM::Parameters())
      : PM::TransformationChecker("DifferentialTransformationChecker", availableParameters(), params), minDiffRotErr(get<float>("minDiffRotErr")),
        minDiffTransErr(get<float>("minDiffTransErr")), smoothLength(get<unsigned>("smoothLength")) {}
  void push(const TP& T) { float q[4]; quatOf(T, q); rot.insert(rot.end(), q, q + 4); tr.push_back(T(0, 3)); tr.push_back(T(1, 3)); tr.push_back(T(2, 3)); }
  void init(const TP& T, bool&) override { rot.clear(); tr.clear(); push(T); }
  void check(const TP& T, bool& iterate) override {
    push(T);
    const size_t cnt = rot.size() / 4, sl = smoothLength;
    float cv0 = 0.f, cv1 = 0.f;
    if (cnt > sl) {
      for (size_t i = cnt - 1; sl > 0 && i >= cnt - sl; --i) {
        cv0 += std::fabs(quatDist(&rot[4 * i], &rot[4 * (i - 1)]));
        const float dx = tr[3 * i] - tr[3 * (i - 1)], dy = tr[3 * i + 1] - tr[3 * (i - 1) + 1], dz = tr[3 * i + 2] - tr[3 * (i - 1) + 2];
        cv1 += std::fabs(std::sqrt(dx * dx + dy * dy + dz * dz));
      }
      cv0 /= (float)sl; cv1 /= (float)sl;
      if (cv0 < minDiffRotErr && cv1 < minDiffTransErr) iterate = false;
    }
    if (cv0 != cv0) throw PM::ConvergenceError("abs rotation norm not a number");
    if (cv1 != cv1) throw PM::ConvergenceError("abs translation norm not a number");
  }
};
struct BoundTransformationChecker : public PM::TransformationChecker {
  static const std::string description() { return "This checker stops the ICP loop with an exception when the transformation values exceed bounds."; }
  static const PM::ParametersDoc availableParameters() {
    return {{"maxRotationNorm", "rotation bound", "1", "0", "inf", &PM::Parametrizable::Comp<float>},
            {"maxTranslationNorm", "translation bound", "1", "0", "inf", &PM::Parametrizable::Comp<float>}};
  }
  const float maxRotationNorm, maxTranslationNorm;
  float q0[4], t0[3];
  explicit BoundTransformationChecker(const PM::Parameters& params = PM::Parameters())
      : PM::TransformationChecker("BoundTransformationChecker", availableParameters(), params), maxRotationNorm(get<float>("maxRotationNorm")),
        maxTranslationNorm(get<float>("maxTranslationNorm")) {}
  void init(const TP& T, bool&) override { quatOf(T, q0); t0[0] = T(0, 3); t0[1] = T(1, 3); t0[2] = T(2, 3); }
  void check(const TP& T, bool&) override {
    float q[4];
    quatOf(T, q);
    const float r = quatDist(q, q0), dx = T(0, 3) - t0[0], dy = T(1, 3) - t0[1], dz = T(2, 3) - t0[2];
    const float t = std::sqrt(dx * dx + dy * dy + dz * dz);
    if (r > maxRotationNorm || t > maxTranslationNorm) {
      std::ostringstream oss;
      oss << "limit out of bounds: rot: " << r << "/" << maxRotationNorm << " tr: " << t << "/" << maxTranslationNorm;
      throw PM::ConvergenceError(oss.str());
    }
  }
};
struct IdentityDataPointsFilter : public PM::DataPointsFilter {
  static const std::string description() { return "Does nothing."; }
  PM::DataPoints filter(const PM::DataPoints& input) override { return input; }
  void inPlaceFilter(PM::DataPoints&) override {}
};
struct NullInspector : public PM::Inspector { static const std::string description() { return "Does nothing."; } };
struct NullLogger : public PM::Logger { static const std::string description() { return "Does not log anything."; } };
}  // namespace b200

// PointMatcher<T>::PointMatcher() (Registry.cpp:61-128): the host-side modules; the CUDA ones come from registerCudaModules()
template <typename T>
PointMatcher<T>::PointMatcher() {
  using namespace b200;
  typedef typename PointMatcher<T>::Transformation Transformation;
  typedef typename PointMatcher<T>::DataPointsFilter DataPointsFilter;
  typedef typename PointMatcher<T>::TransformationChecker TransformationChecker;
  typedef typename PointMatcher<T>::Inspector Inspector;
  typedef typename PointMatcher<T>::Logger Logger;
  ADD_TO_REGISTRAR_NO_PARAM(Transformation, RigidTransformation, RigidTransformation)
  ADD_TO_REGISTRAR_NO_PARAM(DataPointsFilter, IdentityDataPointsFilter, IdentityDataPointsFilter)
  ADD_TO_REGISTRAR(TransformationChecker, CounterTransformationChecker, CounterTransformationChecker)
  ADD_TO_REGISTRAR(TransformationChecker, DifferentialTransformationChecker, DifferentialTransformationChecker)
  ADD_TO_REGISTRAR(TransformationChecker, BoundTransformationChecker, BoundTransformationChecker)
  ADD_TO_REGISTRAR_NO_PARAM(Inspector, NullInspector, NullInspector)
  ADD_TO_REGISTRAR_NO_PARAM(Logger, NullLogger, NullLogger)
}

template <typename T>
const PointMatcher<T>& PointMatcher<T>::get() {   // Registry.cpp:132-148
  static PointMatcher<T> pm;
  static std::once_flag once;
  std::call_once(once, [] { b200::registerCudaModules(const_cast<PointMatcher<T>&>(pm)); });
  return pm;
}

// ---------------------------------------------------------------------------------------------------------------
// ICPChainBase (ICP.cpp:57-232)
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
void PointMatcher<T>::ICPChainBase::cleanup() {   // ICP.cpp:73-86
  transformations.clear();
  readingDataPointsFilters.clear();
  readingStepDataPointsFilters.clear();
  referenceDataPointsFilters.clear();
  matcher.reset();
  outlierFilters.clear();
  errorMinimizer.reset();
  transformationCheckers.clear();
  inspector.reset();
  logger.reset();
}

// ICP.cpp:96-109 with the CUDA twins.  The reference's default chain also has RandomSamplingDataPointsFilter (reading) and
// SamplingSurfaceNormalDataPointsFilter (reference): both are random, out of the hot path's scope and not built here.
template <typename T>
void PointMatcher<T>::ICPChainBase::setDefault() {
  cleanup();
  const PointMatcher& pm = PointMatcher::get();
  transformations.push_back(pm.TransformationRegistrar.create("RigidTransformation"));
  outlierFilters.push_back(pm.OutlierFilterRegistrar.create("CudaTrimmedDistOutlierFilter"));
  matcher = pm.MatcherRegistrar.create("CudaKDTreeMatcher");
  errorMinimizer = pm.ErrorMinimizerRegistrar.create("CudaPointToPlaneErrorMinimizer");
  transformationCheckers.push_back(pm.TransformationCheckerRegistrar.create("CounterTransformationChecker"));
  transformationCheckers.push_back(pm.TransformationCheckerRegistrar.create("DifferentialTransformationChecker"));
  inspector = pm.InspectorRegistrar.create("NullInspector");
}

namespace {
template <typename R>
void createModulesFromRegistrar(const std::string& regName, const YamlNode& doc, const R& registrar,
                                std::vector<std::shared_ptr<typename R::TargetType>>& modules, std::set<std::string>& used) {   // ICP.cpp:185-200
  used.insert(regName);
  const YamlNode* reg = doc.find(regName);
  if (!reg) return;
  if (reg->kind == YamlNode::List) {
    for (const YamlNode& module : reg->list) {
      std::string name;
      Parametrizable::Parameters params;
      getNameParamsFromYAML(module, name, params);
      modules.push_back(registrar.create(name, params));
    }
  } else if (reg->kind == YamlNode::Scalar || (reg->kind == YamlNode::Map && reg->size() > 0)) {
    throw ConfigurationError("YAML: section " + regName + " must be a list of modules");
  }
}
template <typename R>
void createModuleFromRegistrar(const std::string& regName, const YamlNode& doc, const R& registrar,
                               std::shared_ptr<typename R::TargetType>& module, std::set<std::string>& used) {   // ICP.cpp:203-216
  used.insert(regName);
  const YamlNode* reg = doc.find(regName);
  if (reg && reg->kind != YamlNode::Null) {
    std::string name;
    Parametrizable::Parameters params;
    getNameParamsFromYAML(*reg, name, params);
    module = registrar.create(name, params);
  } else {
    module.reset();
  }
}
}  // namespace

template <typename T>
void PointMatcher<T>::ICPChainBase::loadFromYamlText(const std::string& text) {   // ICP.cpp:113-160
  cleanup();
  const YamlNode doc = parseYaml(text);
  std::set<std::string> used;
  const PointMatcher& pm = PointMatcher::get();
  createModuleFromRegistrar("logger", doc, pm.LoggerRegistrar, logger, used);
  createModulesFromRegistrar("readingDataPointsFilters", doc, pm.DataPointsFilterRegistrar, readingDataPointsFilters, used);
  createModulesFromRegistrar("readingStepDataPointsFilters", doc, pm.DataPointsFilterRegistrar, readingStepDataPointsFilters, used);
  createModulesFromRegistrar("referenceDataPointsFilters", doc, pm.DataPointsFilterRegistrar, referenceDataPointsFilters, used);
  createModuleFromRegistrar("matcher", doc, pm.MatcherRegistrar, matcher, used);
  createModulesFromRegistrar("outlierFilters", doc, pm.OutlierFilterRegistrar, outlierFilters, used);
  createModuleFromRegistrar("errorMinimizer", doc, pm.ErrorMinimizerRegistrar, errorMinimizer, used);
  // the Transformation is implicit (ICP.cpp:139-142); only the rigid one exists on this path
  transformations.push_back(pm.TransformationRegistrar.create("RigidTransformation"));
  createModulesFromRegistrar("transformationCheckers", doc, pm.TransformationCheckerRegistrar, transformationCheckers, used);
  createModuleFromRegistrar("inspector", doc, pm.InspectorRegistrar, inspector, used);
  for (auto& kv : doc.map)   // ICP.cpp:152-159
    if (used.find(kv.first) == used.end()) throw InvalidModuleType("Module type " + kv.first + " does not exist");
}

template <typename T>
void PointMatcher<T>::ICPChainBase::loadFromYaml(std::istream& in) {
  std::stringstream ss;
  ss << in.rdbuf();
  loadFromYamlText(ss.str());
}

// ---------------------------------------------------------------------------------------------------------------
// ICP (ICP.cpp:237-468)
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct FastIcp {   // the device-resident ICP of the fast lane
  b200reg_ctx* ctx = nullptr;
  b200reg_icp* icp = nullptr;
  b200reg_icp_config cfg{};
  ~FastIcp() { if (icp) b200reg_icp_destroy(icp); }
};

b200reg_cloud* uploadDataPoints(b200reg_ctx* ctx, const PM::DataPoints& dp) {
  const int64_t n = dp.features.cols();
  std::vector<float> normals;
  const float* np = nullptr;
  if (dp.descriptorExists("normals") && n > 0) {
    const int row = dp.getDescriptorStartingRow("normals");
    normals.resize(3 * (size_t)n);
    for (int64_t c = 0; c < n; ++c) for (int r = 0; r < 3; ++r) normals[3 * c + r] = dp.descriptors(row + r, c);
    np = normals.data();
  }
  b200reg_cloud* cloud = nullptr;
  const int32_t st = b200reg_cloud_upload_f32(ctx, dp.features.data(), (int32_t)dp.features.rows(), np, n, &cloud);
  if (st != B200REG_OK) b200::throwStatus(ctx, st);
  return cloud;
}
}  // namespace

template <typename T>
bool PointMatcher<T>::ICP::fastLaneEligible() const {
  using namespace b200;
  if (!dynamic_cast<CudaKDTreeMatcher*>(this->matcher.get())) return false;
  if (!dynamic_cast<CudaPointToPlaneErrorMinimizer*>(this->errorMinimizer.get())) return false;
  if (dynamic_cast<CudaPointToPlaneWithCovErrorMinimizer*>(this->errorMinimizer.get())) return false;   // its covariance is a per-module product
  if (!this->readingStepDataPointsFilters.empty()) return false;
  int nTrim = 0, nAngle = 0, nCnt = 0, nDiff = 0, nBound = 0;
  for (auto& f : this->outlierFilters) {
    if (dynamic_cast<CudaTrimmedDistOutlierFilter*>(f.get())) ++nTrim;
    else if (dynamic_cast<CudaSurfaceNormalOutlierFilter*>(f.get())) ++nAngle;
    else return false;
  }
  for (auto& c : this->transformationCheckers) {
    if (dynamic_cast<CounterTransformationChecker*>(c.get())) ++nCnt;
    else if (dynamic_cast<DifferentialTransformationChecker*>(c.get())) ++nDiff;
    else if (dynamic_cast<BoundTransformationChecker*>(c.get())) ++nBound;
    else return false;
  }
  return nTrim <= 1 && nAngle <= 1 && nCnt <= 1 && nDiff <= 1 && nBound <= 1 && this->transformations.size() == 1;
}

template <typename T>
typename PointMatcher<T>::TransformationParameters PointMatcher<T>::ICP::operator()(const DataPoints& readingIn, const DataPoints& referenceIn) {
  return this->compute(readingIn, referenceIn, Matrix::Identity((int)readingIn.features.rows()));   // ICP.cpp:237-244
}
template <typename T>
typename PointMatcher<T>::TransformationParameters PointMatcher<T>::ICP::operator()(const DataPoints& readingIn, const DataPoints& referenceIn,
                                                                                    const TransformationParameters& T_init) {
  return this->compute(readingIn, referenceIn, T_init);
}

template <typename T>
typename PointMatcher<T>::TransformationParameters PointMatcher<T>::ICP::compute(const DataPoints& readingIn, const DataPoints& referenceIn,
                                                                                  const TransformationParameters& T_refIn_readIn,
                                                                                  const bool initializeMatcherWithInputReference) {
  if (!this->matcher) throw std::runtime_error("You must setup a matcher before running ICP");               // ICP.cpp:265-270
  if (!this->errorMinimizer) throw std::runtime_error("You must setup an error minimizer before running ICP");
  if (!this->inspector) throw std::runtime_error("You must setup an inspector before running ICP");
  if (initializeMatcherWithInputReference || !matcherIsInitialized) initReference(referenceIn);                // :277-279
  this->inspector->addStat("ReferenceInPointCount", (double)referenceIn.features.cols());
  this->inspector->addStat("ReferencePointCount", (double)referenceFiltered.features.cols());
  this->prefilteredReferencePtsCount = (unsigned)referenceFiltered.features.cols();
  return computeWithTransformedReference(readingIn, referenceFiltered, T_refIn_refMean, T_refIn_readIn);
}

template <typename T>
bool PointMatcher<T>::ICP::initReference(const DataPoints& referenceIn) {   // ICP.cpp:293-328
  const int64_t nbPtsReference = referenceIn.features.cols();
  if (nbPtsReference == 0) { matcherIsInitialized = false; return false; }
  const int dim = (int)referenceIn.features.rows();
  referenceFiltered = referenceIn;
  T_refIn_refMean = Matrix::Identity(dim);
  for (auto& f : this->referenceDataPointsFilters) { f->init(); f->inPlaceFilter(referenceFiltered); }
  const int64_t m = referenceFiltered.features.cols();
  if (fastLaneEligible()) {
    using namespace b200;
    auto* km = static_cast<CudaKDTreeMatcher*>(this->matcher.get());
    auto fi = std::make_shared<FastIcp>();
    fi->ctx = contextFor(km->device);
    b200reg_icp_config& c = fi->cfg;
    c.knn = km->knn; c.max_dist = km->maxDist; c.epsilon = km->epsilon; c.cell = km->cell;
    c.trim_ratio = -1.f; c.max_angle = -1.f; c.max_iter = -1; c.smooth_length = -1; c.bound_rot = -1.f; c.bound_trans = -1.f;
    for (auto& f : this->outlierFilters) {
      if (auto* t = dynamic_cast<CudaTrimmedDistOutlierFilter*>(f.get())) c.trim_ratio = t->ratio;
      if (auto* a = dynamic_cast<CudaSurfaceNormalOutlierFilter*>(f.get())) c.max_angle = a->maxAngle;
    }
    auto* em = static_cast<CudaPointToPlaneErrorMinimizer*>(this->errorMinimizer.get());
    c.force2D = em->force2D; c.force4DOF = em->force4DOF;
    int slot = 0;   // the device loop evaluates the checkers in the chain's order (ICP.cpp:437-446)
    c.checker_order[0] = c.checker_order[1] = c.checker_order[2] = 0;
    for (auto& ch : this->transformationCheckers) {
      if (auto* x = dynamic_cast<CounterTransformationChecker*>(ch.get())) { c.max_iter = (int)std::min<unsigned>(x->maxIterationCount, 2147483647u); c.checker_order[slot++] = B200REG_CHECKER_COUNTER; }
      if (auto* x = dynamic_cast<DifferentialTransformationChecker*>(ch.get())) { c.min_diff_rot = x->minDiffRotErr; c.min_diff_trans = x->minDiffTransErr; c.smooth_length = (int)x->smoothLength; c.checker_order[slot++] = B200REG_CHECKER_DIFFERENTIAL; }
      if (auto* x = dynamic_cast<BoundTransformationChecker*>(ch.get())) { c.bound_rot = x->maxRotationNorm; c.bound_trans = x->maxTranslationNorm; c.checker_order[slot++] = B200REG_CHECKER_BOUND; }
    }
    int32_t st = b200reg_icp_create(fi->ctx, &c, &fi->icp);
    if (st != B200REG_OK) throwStatus(fi->ctx, st);
    b200reg_cloud* cloud = uploadDataPoints(fi->ctx, referenceFiltered);
    st = b200reg_icp_set_reference(fi->icp, cloud);
    b200reg_cloud_free(fi->ctx, cloud);
    if (st != B200REG_OK) throwStatus(fi->ctx, st);
    float mean[3];
    b200reg_icp_reference_mean(fi->icp, mean);
    for (int d = 0; d < 3; ++d) T_refIn_refMean(d, dim - 1) = mean[d];
    fastIcp = fi;
    this->matcher->resetVisitCount();
    matcherIsInitialized = true;
    return true;
  }
  fastIcp.reset();
  // meanReference = features.rowwise().mean() (:313) — contract: fp64 sum rounded once to fp32
  double sum[3] = {0, 0, 0};
  for (int64_t c = 0; c < m; ++c) for (int d = 0; d < dim - 1; ++d) sum[d] += referenceFiltered.features(d, c);
  float mean[3] = {0, 0, 0};
  for (int d = 0; d < dim - 1; ++d) { mean[d] = (float)(sum[d] / (double)m); T_refIn_refMean(d, dim - 1) = mean[d]; }
  for (int64_t c = 0; c < m; ++c) for (int d = 0; d < dim - 1; ++d) referenceFiltered.features(d, c) -= mean[d];   // :320
  this->matcher->resetVisitCount();
  this->matcher->init(referenceFiltered);   // :324
  matcherIsInitialized = true;
  return true;
}

template <typename T>
typename PointMatcher<T>::TransformationParameters PointMatcher<T>::ICP::computeWithTransformedReference(
    const DataPoints& readingIn, const DataPoints& reference, const TransformationParameters& T_refIn_refMeanArg,
    const TransformationParameters& T_refIn_readIn) {
  const int dim = (int)reference.features.rows();
  if (T_refIn_readIn.cols() != T_refIn_readIn.rows()) throw std::runtime_error("The initial transformation matrix must be squared.");
  if (dim != T_refIn_readIn.cols())
    throw std::runtime_error("The shape of initial transformation matrix must be NxN. Where N is the number of rows in the read/reference scans.");
  DataPoints reading(readingIn);   // :352
  for (auto& f : this->readingDataPointsFilters) { f->init(); f->inPlaceFilter(reading); }
  if (reading.getNbPoints() == 0) throw std::runtime_error("The reading point cloud is empty.");   // :357-359
  this->inspector->addStat("ReadingInPointCount", (double)readingIn.features.cols());
  this->inspector->addStat("ReadingPointCount", (double)reading.features.cols());
  this->prefilteredReadingPtsCount = (unsigned)reading.features.cols();
  this->maxNumIterationsReached = false;

  if (fastIcp) {   // whole iterate loop on the device
    FastIcp* fi = static_cast<FastIcp*>(fastIcp.get());
    b200reg_cloud* cloud = uploadDataPoints(fi->ctx, reading);
    float Ti[16], To[16];
    toRowMajor(T_refIn_readIn, Ti);
    b200reg_icp_stats st;
    const int32_t code = b200reg_icp_register(fi->icp, cloud, Ti, To, &st);
    b200reg_cloud_free(fi->ctx, cloud);
    lastFastLane = true;
    lastIterations = st.iterations;
    this->maxNumIterationsReached = st.max_iterations_reached != 0;
    this->errorMinimizer->pointUsedRatio = st.point_used_ratio;
    this->errorMinimizer->weightedPointUsedRatio = st.overlap_ratio;
    this->inspector->addStat("IterationsCount", st.iterations);
    this->inspector->addStat("PointCountTouched", (double)st.points_touched);
    this->inspector->addStat("OverlapRatio", st.overlap_ratio);
    this->inspector->addStat("ConvergenceDuration", st.convergence_s);
    if (code != B200REG_OK) b200::throwStatus(fi->ctx, code);
    readingFiltered = reading;
    return fromRowMajor(To);
  }
  lastFastLane = false;

  // meanReading := 0 in this fork (:363-364) => T_readIn_readMean = I
  TP T_refMean_refIn = Matrix::Identity(dim);
  for (int d = 0; d < dim - 1; ++d) T_refMean_refIn(d, dim - 1) = -T_refIn_refMeanArg(d, dim - 1);
  const TP T_refMean_readMean = mul4(T_refMean_refIn, T_refIn_readIn);   // :373
  for (auto& t : this->transformations) t->inPlaceCompute(T_refMean_readMean, reading);   // :375
  for (auto& f : this->readingStepDataPointsFilters) f->init();
  TP T_iter = Matrix::Identity(dim);
  bool iterate = true;
  for (auto& c : this->transformationCheckers) c->init(T_iter, iterate);   // :386
  size_t iterationCount = 0;
  while (iterate) {   // :399-448
    DataPoints stepReading(reading);
    for (auto& f : this->readingStepDataPointsFilters) f->inPlaceFilter(stepReading);
    for (auto& t : this->transformations) t->inPlaceCompute(T_iter, stepReading);
    const Matches matches = this->matcher->findClosests(stepReading);
    OutlierWeights w;
    if (this->outlierFilters.empty()) {   // OutlierFilter.cpp:70-85
      w = OutlierWeights(matches.dists.rows(), matches.dists.cols());
      for (size_t i = 0; i < w.d.size(); ++i) w.d[i] = (matches.dists.d[i] == std::numeric_limits<T>::infinity()) ? T(0) : T(1);
    } else {
      w = this->outlierFilters[0]->compute(stepReading, reference, matches);
      for (size_t f = 1; f < this->outlierFilters.size(); ++f) {
        const OutlierWeights w2 = this->outlierFilters[f]->compute(stepReading, reference, matches);
        for (size_t i = 0; i < w.d.size(); ++i) w.d[i] = w.d[i] * w2.d[i];
      }
    }
    T_iter = mul4(this->errorMinimizer->compute(stepReading, reference, w, matches), T_iter);   // :433
    try {
      for (auto& c : this->transformationCheckers) c->check(T_iter, iterate);
    } catch (const b200::CounterTransformationChecker::MaxNumIterationsReached&) {   // :441-445
      iterate = false;
      this->maxNumIterationsReached = true;
    }
    ++iterationCount;
  }
  lastIterations = (int)iterationCount;
  this->inspector->addStat("IterationsCount", (double)iterationCount);
  this->inspector->addStat("PointCountTouched", (double)this->matcher->getVisitCount());
  this->matcher->resetVisitCount();
  this->inspector->addStat("OverlapRatio", this->errorMinimizer->getWeightedPointUsedRatio());
  readingFiltered = reading;
  return mul4(T_refIn_refMeanArg, mul4(T_iter, T_refMean_readMean));   // :459-465
}

template struct PointMatcher<float>;
