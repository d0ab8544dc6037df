// host/pm_compat.h — minimal, header-declared stand-ins for the parts of the libpointmatcher API the CUDA modules touch.
//
// WHY THIS EXISTS: the reference's host language is C++ and its toolchain (g++) is present, but libpointmatcher's
// mandatory dependencies (Eigen, boost, yaml-cpp, libnabo) are not installed and cannot be (no network), so
// `#include "pointmatcher/PointMatcher.h"` cannot be preprocessed here.  This header mirrors — same names, same argument
// meaning, same error behaviour — the registrable-module surface of
//   libpointmatcher/pointmatcher/PointMatcher.h:229-848, Parametrizable.h:98-175, Registrar.h:58-213, ICP.cpp:96-468
// so that host/cuda_modules.{h,cpp} read like modules written against the real library.  INTEGRATION.md shows the
// handful of lines that change when compiling against real libpointmatcher.  Column-major matrices like Eigen's.
#pragma once
#include <cmath>
#include <cstdint>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace PointMatcherSupport {

// ---- lexical_cast_scalar_to_string (Parametrizable.h:53-64): "inf", "-inf", "nan" aware ----------------------------
struct InvalidParameter : std::runtime_error { explicit InvalidParameter(const std::string& r) : std::runtime_error(r) {} };
struct InvalidElement : std::runtime_error { explicit InvalidElement(const std::string& r) : std::runtime_error(r) {} };
struct ConfigurationError : std::runtime_error { explicit ConfigurationError(const std::string& r) : std::runtime_error(r) {} };

template <typename Target>
inline Target lexical_cast(const std::string& arg) {
  if (std::numeric_limits<Target>::has_infinity) {
    if (arg == "inf" || arg == "+inf") return std::numeric_limits<Target>::infinity();
    if (arg == "-inf") return -std::numeric_limits<Target>::infinity();
    if (arg == "nan") return std::numeric_limits<Target>::quiet_NaN();
  }
  std::istringstream is(arg);
  Target t{};
  is >> t;
  if (is.fail() || !(is >> std::ws).eof()) throw InvalidParameter("bad lexical cast: '" + arg + "'");
  return t;
}
template <>
inline std::string lexical_cast<std::string>(const std::string& arg) { return arg; }
template <>
inline bool lexical_cast<bool>(const std::string& arg) {
  if (arg == "1" || arg == "true") return true;
  if (arg == "0" || arg == "false") return false;
  throw InvalidParameter("bad lexical cast to bool: '" + arg + "'");
}
template <typename S>
inline std::string toParam(const S& value) { std::ostringstream os; os << value; return os.str(); }

// ---- Parametrizable (Parametrizable.h:98-175, Parametrizable.cpp:170-206) ------------------------------------------
struct Parametrizable {
  typedef PointMatcherSupport::InvalidParameter InvalidParameter;
  typedef bool (*LexicalComparison)(std::string a, std::string b);
  template <typename S>
  static bool Comp(std::string a, std::string b) { return lexical_cast<S>(a) < lexical_cast<S>(b); }
  struct ParameterDoc {
    std::string name, doc, defaultValue, minValue, maxValue;
    LexicalComparison comp;
    ParameterDoc(const std::string& n, const std::string& d, const std::string& def, const std::string& mn, const std::string& mx, LexicalComparison c)
        : name(n), doc(d), defaultValue(def), minValue(mn), maxValue(mx), comp(c) {}
    ParameterDoc(const std::string& n, const std::string& d, const std::string& def) : name(n), doc(d), defaultValue(def), comp(nullptr) {}
  };
  typedef std::vector<ParameterDoc> ParametersDoc;
  typedef std::string Parameter;
  typedef std::map<std::string, Parameter> Parameters;
  typedef std::set<std::string> ParametersUsed;

  const std::string className;
  const ParametersDoc parametersDoc;
  Parameters parameters;
  ParametersUsed parametersUsed;

  Parametrizable() : className("unknown") {}
  Parametrizable(const std::string& className, const ParametersDoc paramsDoc, const Parameters& params);
  virtual ~Parametrizable() {}
  std::string getParamValueString(const std::string& paramName);
  template <typename S>
  S get(const std::string& paramName) { return lexical_cast<S>(getParamValueString(paramName)); }
};

// ---- Registrar (Registrar.h:58-201) ------------------------------------------------------------------------------------
template <typename Interface>
struct Registrar {
  typedef Interface TargetType;
  struct ClassDescriptor {
    virtual ~ClassDescriptor() {}
    virtual std::shared_ptr<Interface> createInstance(const std::string& className, const Parametrizable::Parameters& params) const = 0;
    virtual const std::string description() const = 0;
    virtual const Parametrizable::ParametersDoc availableParameters() const = 0;
  };
  template <typename C>
  struct GenericClassDescriptor : public ClassDescriptor {
    std::shared_ptr<Interface> createInstance(const std::string& className, const Parametrizable::Parameters& params) const override {
      std::shared_ptr<C> instance = std::make_shared<C>(params);
      for (auto it = params.begin(); it != params.end(); ++it)   // Registrar.h:86-93
        if (instance->parametersUsed.find(it->first) == instance->parametersUsed.end())
          throw Parametrizable::InvalidParameter("Parameter " + it->first + " for module " + className + " was set but is not used");
      return instance;
    }
    const std::string description() const override { return C::description(); }
    const Parametrizable::ParametersDoc availableParameters() const override { return C::availableParameters(); }
  };
  template <typename C>
  struct GenericClassDescriptorNoParam : public ClassDescriptor {
    std::shared_ptr<Interface> createInstance(const std::string& className, const Parametrizable::Parameters& params) const override {
      for (auto it = params.begin(); it != params.end(); ++it)
        throw Parametrizable::InvalidParameter("Parameter " + it->first + " was set but module " + className + " dos not use any parameter");
      return std::make_shared<C>();
    }
    const std::string description() const override { return C::description(); }
    const Parametrizable::ParametersDoc availableParameters() const override { return Parametrizable::ParametersDoc(); }
  };

  std::map<std::string, ClassDescriptor*> classes;
  ~Registrar() { for (auto& kv : owned) delete kv; }
  void reg(const std::string& name, ClassDescriptor* descriptor) { classes[name] = descriptor; owned.insert(descriptor); }   // Registrar.h:141-144
  const ClassDescriptor* getDescriptor(const std::string& name) const {
    auto it = classes.find(name);
    if (it == classes.end()) throw InvalidElement("Trying to instanciate unknown element " + name + " from registrar");   // :146-158
    return it->second;
  }
  std::shared_ptr<Interface> create(const std::string& name, const Parametrizable::Parameters& params = Parametrizable::Parameters()) const {
    return getDescriptor(name)->createInstance(name, params);
  }
  const std::string getDescription(const std::string& name) const { return getDescriptor(name)->description(); }
  std::vector<std::string> names() const { std::vector<std::string> r; for (auto& kv : classes) r.push_back(kv.first); return r; }
 private:
  std::set<ClassDescriptor*> owned;
};

#define ADD_TO_REGISTRAR(name, elementName, element) \
  name##Registrar.reg(#elementName, new typename PointMatcherSupport::Registrar<name>::template GenericClassDescriptor<element>());
#define ADD_TO_REGISTRAR_NO_PARAM(name, elementName, element) \
  name##Registrar.reg(#elementName, new typename PointMatcherSupport::Registrar<name>::template GenericClassDescriptorNoParam<element>());

// ---- a YAML-subset node (the ICP chain files only use maps, lists, scalars and comments); parser in pm_compat.cpp ----
struct YamlNode {
  enum Kind { Null, Scalar, Map, List } kind = Null;
  std::string scalar;
  std::vector<std::pair<std::string, YamlNode>> map;   // insertion order kept
  std::vector<YamlNode> list;
  const YamlNode* find(const std::string& key) const { for (auto& kv : map) if (kv.first == key) return &kv.second; return nullptr; }
  size_t size() const { return kind == Map ? map.size() : kind == List ? list.size() : 0; }
};
YamlNode parseYaml(const std::string& text);
// getNameParamsFromYAML (Registrar.cpp:7-26)
void getNameParamsFromYAML(const YamlNode& module, std::string& name, Parametrizable::Parameters& params);

}  // namespace PointMatcherSupport

// ============================================================================================================================
// PointMatcher<T> — only T = float is instantiated (usings.h:13 fixes float on the hot path)
// ============================================================================================================================
template <typename T>
struct PointMatcher {
  typedef PointMatcherSupport::Parametrizable Parametrizable;
  typedef Parametrizable::Parameters Parameters;
  typedef Parametrizable::ParametersDoc ParametersDoc;
  typedef PointMatcherSupport::InvalidParameter InvalidParameter;

  struct ConvergenceError : std::runtime_error { explicit ConvergenceError(const std::string& r) : std::runtime_error(r) {} };   // PointMatcher.h:100-103
  struct TransformationError : std::runtime_error { explicit TransformationError(const std::string& r) : std::runtime_error(r) {} };
  struct InvalidModuleType : std::runtime_error { explicit InvalidModuleType(const std::string& r) : std::runtime_error(r) {} };

  // column-major dense matrix with Eigen's memory layout
  template <typename S>
  struct Mat {
    int64_t r = 0, c = 0;
    std::vector<S> d;
    Mat() {}
    Mat(int64_t rows, int64_t cols, S v = S()) : r(rows), c(cols), d((size_t)rows * cols, v) {}
    int64_t rows() const { return r; }
    int64_t cols() const { return c; }
    S& operator()(int64_t i, int64_t j) { return d[(size_t)j * r + i]; }
    const S& operator()(int64_t i, int64_t j) const { return d[(size_t)j * r + i]; }
    S* data() { return d.data(); }
    const S* data() const { return d.data(); }
    static Mat Identity(int n) { Mat m(n, n, S(0)); for (int i = 0; i < n; ++i) m(i, i) = S(1); return m; }
    void conservativeResizeCols(int64_t cols) { d.resize((size_t)r * cols); c = cols; }
  };
  typedef Mat<T> Matrix;
  typedef Mat<int> IntMatrix;
  typedef Mat<int64_t> Int64Matrix;
  typedef Matrix TransformationParameters;   // 4x4
  typedef Matrix OutlierWeights;             // PointMatcher.h:470

  // ---- DataPoints (PointMatcher.h:229-431) ----
  struct DataPoints {
    struct InvalidField : std::runtime_error { explicit InvalidField(const std::string& r) : std::runtime_error(r) {} };
    struct Label { std::string text; size_t span; Label(const std::string& t = "", size_t s = 0) : text(t), span(s) {} };
    typedef std::vector<Label> Labels;
    Matrix features;        // (dim+1) x N, homogeneous
    Labels featureLabels;
    Matrix descriptors;     // D x N
    Labels descriptorLabels;
    Int64Matrix times;
    Labels timeLabels;
    int64_t getNbPoints() const { return features.cols(); }
    bool descriptorExists(const std::string& name) const { for (auto& l : descriptorLabels) if (l.text == name) return true; return false; }
    // row offset of a descriptor, or throws InvalidField (getDescriptorViewByName semantics)
    int getDescriptorStartingRow(const std::string& name, size_t* span = nullptr) const;
    // DataPoints::allocateDescriptor semantics (DataPoints.cpp:988-1027): reuse same-span label, InvalidField on span mismatch, else append rows
    int allocateDescriptor(const std::string& name, size_t span);
    void conservativeResize(int64_t n);
  };

  // ---- Matches (PointMatcher.h:444-464) ----
  struct Matches {
    typedef Matrix Dists;
    typedef IntMatrix Ids;
    static constexpr int InvalidId = -1;
    static T InvalidDist() { return std::numeric_limits<T>::infinity(); }
    Dists dists;   // knn x N, SQUARED distances
    Ids ids;       // knn x N
    Matches() {}
    Matches(int knn, int64_t n) : dists(knn, n), ids(knn, n) {}
    T getDistsQuantile(const T quantile) const;   // Matches.cpp:61-87 (host restatement for CPU-side users of the shim)
  };

  // ---- module interfaces (PointMatcher.h:477-718) ----
  struct Transformation : public Parametrizable {
    Transformation() {}
    Transformation(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~Transformation() {}
    virtual DataPoints compute(const DataPoints& input, const TransformationParameters& parameters) const = 0;
    virtual void inPlaceCompute(const TransformationParameters& parameters, DataPoints& cloud) const = 0;
    virtual bool checkParameters(const TransformationParameters& parameters) const = 0;
  };
  struct DataPointsFilter : public Parametrizable {
    DataPointsFilter() {}
    DataPointsFilter(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~DataPointsFilter() {}
    virtual void init() {}
    virtual DataPoints filter(const DataPoints& input) = 0;
    virtual void inPlaceFilter(DataPoints& cloud) = 0;
  };
  struct Matcher : public Parametrizable {
    unsigned long visitCounter = 0;
    Matcher() {}
    Matcher(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~Matcher() {}
    void resetVisitCount() { visitCounter = 0; }
    unsigned long getVisitCount() const { return visitCounter; }
    virtual void init(const DataPoints& filteredReference) = 0;
    virtual Matches findClosests(const DataPoints& filteredReading) = 0;
  };
  struct OutlierFilter : public Parametrizable {
    OutlierFilter() {}
    OutlierFilter(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~OutlierFilter() {}
    virtual OutlierWeights compute(const DataPoints& filteredReading, const DataPoints& filteredReference, const Matches& input) = 0;
  };
  struct ErrorMinimizer : public Parametrizable {
    T pointUsedRatio = -1, weightedPointUsedRatio = -1;
    ErrorMinimizer() {}
    ErrorMinimizer(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~ErrorMinimizer() {}
    // the 4-argument form is virtual in the reference too (PointMatcher.h:634-636): overriding it is the sanctioned way to skip ErrorElements
    virtual TransformationParameters compute(const DataPoints& filteredReading, const DataPoints& filteredReference,
                                             const OutlierWeights& outlierWeights, const Matches& matches) = 0;
    T getPointUsedRatio() const { return pointUsedRatio; }
    T getWeightedPointUsedRatio() const { return weightedPointUsedRatio; }
    virtual T getOverlap() const { return weightedPointUsedRatio; }   // ErrorMinimizer.cpp:258-262; PointToPlane.cpp:477-481 takes the same value without sensor-noise descriptors
    virtual T getResidualError(const DataPoints&, const DataPoints&, const OutlierWeights&, const Matches&) const {   // ErrorMinimizer.cpp:274-278
      return std::numeric_limits<T>::max();
    }
    virtual Matrix getCovariance() const { return Matrix(6, 6, T(0)); }   // ErrorMinimizer.cpp:264-270: zero unless a child redefines it
  };
  struct TransformationChecker : public Parametrizable {
    TransformationChecker() {}
    TransformationChecker(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    virtual ~TransformationChecker() {}
    virtual void init(const TransformationParameters& parameters, bool& iterate) = 0;
    virtual void check(const TransformationParameters& parameters, bool& iterate) = 0;
  };
  struct Inspector : public Parametrizable {
    Inspector() {}
    Inspector(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
    std::map<std::string, double> stats;   // the reference's stat names (ICP.cpp:282-284,391-393,452-456)
    virtual void addStat(const std::string& name, double data) { stats[name] = data; }
  };
  struct Logger : public Parametrizable {
    Logger() {}
    Logger(const std::string& cn, const ParametersDoc pd, const Parameters& p) : Parametrizable(cn, pd, p) {}
  };

  // ---- registrars (DEF_REGISTRAR, PointMatcher.h:505,539,564,596,648) ----
  PointMatcherSupport::Registrar<Transformation> TransformationRegistrar;
  PointMatcherSupport::Registrar<DataPointsFilter> DataPointsFilterRegistrar;
  PointMatcherSupport::Registrar<Matcher> MatcherRegistrar;
  PointMatcherSupport::Registrar<OutlierFilter> OutlierFilterRegistrar;
  PointMatcherSupport::Registrar<ErrorMinimizer> ErrorMinimizerRegistrar;
  PointMatcherSupport::Registrar<TransformationChecker> TransformationCheckerRegistrar;
  PointMatcherSupport::Registrar<Inspector> InspectorRegistrar;
  PointMatcherSupport::Registrar<Logger> LoggerRegistrar;

  PointMatcher();                       // registers the host-side modules (Registry.cpp:61-128 analogue)
  static const PointMatcher& get();     // Registry.cpp:132-148

  // ---- ICP chain (PointMatcher.h:729-848, ICP.cpp) ----
  struct ICPChainBase {
    std::vector<std::shared_ptr<DataPointsFilter>> readingDataPointsFilters, readingStepDataPointsFilters, referenceDataPointsFilters;
    std::vector<std::shared_ptr<Transformation>> transformations;
    std::shared_ptr<Matcher> matcher;
    std::vector<std::shared_ptr<OutlierFilter>> outlierFilters;
    std::shared_ptr<ErrorMinimizer> errorMinimizer;
    std::vector<std::shared_ptr<TransformationChecker>> transformationCheckers;
    std::shared_ptr<Inspector> inspector;
    std::shared_ptr<Logger> logger;
    virtual ~ICPChainBase() {}
    virtual void setDefault();                         // ICP.cpp:96-109 with the CUDA twins (the two random samplers are out of scope)
    void loadFromYaml(std::istream& in);               // ICP.cpp:113-160
    void loadFromYamlText(const std::string& text);
    bool getMaxNumIterationsReached() const { return maxNumIterationsReached; }
    unsigned getPrefilteredReadingPtsCount() const { return prefilteredReadingPtsCount; }
    unsigned getPrefilteredReferencePtsCount() const { return prefilteredReferencePtsCount; }
   protected:
    unsigned prefilteredReadingPtsCount = 0, prefilteredReferencePtsCount = 0;
    bool maxNumIterationsReached = false;
    void cleanup();
  };

  struct ICP : ICPChainBase {
    TransformationParameters operator()(const DataPoints& readingIn, const DataPoints& referenceIn);                                  // ICP.cpp:237-244
    TransformationParameters operator()(const DataPoints& readingIn, const DataPoints& referenceIn, const TransformationParameters& T_init);
    TransformationParameters compute(const DataPoints& readingIn, const DataPoints& referenceIn, const TransformationParameters& T_refIn_readIn,
                                     const bool initializeMatcherWithInputReference = true);                                           // ICP.cpp:258-290
    bool initReference(const DataPoints& referenceIn);                                                                                 // ICP.cpp:293-328
    const DataPoints& getReadingFiltered() const { return readingFiltered; }
    // true when the last compute() ran entirely on the device through b200reg_icp_register (all modules CUDA)
    bool usedFastLane() const { return lastFastLane; }
    int getIterationCount() const { return lastIterations; }
   protected:
    TransformationParameters computeWithTransformedReference(const DataPoints& readingIn, const DataPoints& reference,
                                                             const TransformationParameters& T_refIn_refMean,
                                                             const TransformationParameters& T_refIn_readIn);                         // ICP.cpp:332-468
    bool fastLaneEligible() const;
    bool matcherIsInitialized = false;
    DataPoints referenceFiltered, readingFiltered;
    TransformationParameters T_refIn_refMean;
    bool lastFastLane = false;
    int lastIterations = 0;
    std::shared_ptr<void> fastIcp;   // b200reg_icp handle + its context (see pm_compat.cpp)
  };
};

typedef PointMatcher<float> PM;
