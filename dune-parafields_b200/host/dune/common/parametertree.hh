// Minimal stand-in for Dune::ParameterTree (dune-common is not available in
// this environment).  Reproduces the documented public
// behaviour the random field code relies on: hierarchical string storage
// addressed by dotted keys, typed get<T>(key[, default]) with stream parsing,
// bool accepting yes/no/true/false, exact-arity std::array parsing (this is
// what couples the arity of stochastic.corrLength to stochastic.anisotropy),
// RangeError on missing keys, report() as ini text.
#ifndef PF_STANDIN_DUNE_COMMON_PARAMETERTREE_HH
#define PF_STANDIN_DUNE_COMMON_PARAMETERTREE_HH

#include <algorithm>
#include <array>
#include <cctype>
#include <cstddef>
#include <iostream>
#include <iterator>
#include <map>
#include <sstream>
#include <string>
#include <typeinfo>
#include <vector>

#include <dune/common/exceptions.hh>
#include <dune/common/fvector.hh>

namespace Dune {

class ParameterTree
{
  template<class T, class = void>
  struct Parser;

public:
  typedef std::vector<std::string> KeyVector;

  ParameterTree() = default;

  bool hasKey(const std::string& key) const
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos) {
      auto it = _subs.find(key.substr(0, dot));
      if (it == _subs.end())
        return false;
      return it->second.hasKey(key.substr(dot + 1));
    }
    return _values.count(key) != 0;
  }

  bool hasSub(const std::string& key) const
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos) {
      auto it = _subs.find(key.substr(0, dot));
      if (it == _subs.end())
        return false;
      return it->second.hasSub(key.substr(dot + 1));
    }
    return _subs.count(key) != 0;
  }

  std::string& operator[](const std::string& key)
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos)
      return sub(key.substr(0, dot))[key.substr(dot + 1)];
    if (!_values.count(key)) {
      _values[key] = "";
      _valueKeys.push_back(key);
    }
    return _values[key];
  }

  const std::string& operator[](const std::string& key) const
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos)
      return sub(key.substr(0, dot))[key.substr(dot + 1)];
    auto it = _values.find(key);
    if (it == _values.end())
      DUNE_THROW(RangeError, "Key '" << key
                                     << "' not found in ParameterTree (prefix "
                                     << _prefix << ")");
    return it->second;
  }

  ParameterTree& sub(const std::string& key)
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos)
      return sub(key.substr(0, dot)).sub(key.substr(dot + 1));
    if (!_subs.count(key)) {
      _subKeys.push_back(key);
      _subs[key]._prefix = _prefix + key + ".";
    }
    return _subs[key];
  }

  const ParameterTree& sub(const std::string& key,
                           bool fail_if_missing = false) const
  {
    const auto dot = key.find('.');
    if (dot != std::string::npos)
      return sub(key.substr(0, dot), fail_if_missing)
        .sub(key.substr(dot + 1), fail_if_missing);
    auto it = _subs.find(key);
    if (it == _subs.end()) {
      if (_values.count(key))
        DUNE_THROW(RangeError,
                   "'" << key << "' is a value key, not a sub tree (prefix "
                       << _prefix << ")");
      if (fail_if_missing)
        DUNE_THROW(RangeError, "SubTree '"
                                 << key
                                 << "' not found in ParameterTree (prefix "
                                 << _prefix << ")");
      return empty();
    }
    return it->second;
  }

  const KeyVector& getValueKeys() const { return _valueKeys; }
  const KeyVector& getSubKeys() const { return _subKeys; }

  // ---- typed access --------------------------------------------------------
  std::string get(const std::string& key, const std::string& defaultValue) const
  {
    return hasKey(key) ? (*this)[key] : defaultValue;
  }
  std::string get(const std::string& key, const char* defaultValue) const
  {
    return hasKey(key) ? (*this)[key] : std::string(defaultValue);
  }
  template<typename T>
  T get(const std::string& key, const T& defaultValue) const
  {
    return hasKey(key) ? get<T>(key) : defaultValue;
  }
  template<class T>
  T get(const std::string& key) const
  {
    if (!hasKey(key))
      DUNE_THROW(RangeError, "Key '" << key
                                     << "' not found in ParameterTree (prefix "
                                     << _prefix << ")");
    try {
      return Parser<T>::parse((*this)[key]);
    } catch (const RangeError& e) {
      DUNE_THROW(RangeError, "Cannot parse value \""
                               << (*this)[key] << "\" for key \"" << _prefix
                               << "." << key << "\": " << e.what());
    }
  }

  // ini-style dump (used for the .field / .fieldList files)
  void report(std::ostream& stream = std::cout,
              const std::string& prefix = "") const
  {
    for (const auto& k : _valueKeys)
      stream << k << " = \"" << _values.at(k) << "\"" << std::endl;
    for (const auto& k : _subKeys) {
      stream << "[ " << prefix + k << " ]" << std::endl;
      _subs.at(k).report(stream, prefix + k + ".");
    }
  }

private:
  static const ParameterTree& empty()
  {
    static const ParameterTree e;
    return e;
  }

  static std::string ltrim(const std::string& s)
  {
    std::size_t i = 0;
    while (i < s.size() && std::isspace((unsigned char)s[i]))
      i++;
    return s.substr(i);
  }
  static std::string rtrim(const std::string& s)
  {
    std::size_t n = s.size();
    while (n > 0 && std::isspace((unsigned char)s[n - 1]))
      n--;
    return s.substr(0, n);
  }
  static std::vector<std::string> split(const std::string& s)
  {
    std::vector<std::string> out;
    std::istringstream in(s);
    std::string tok;
    while (in >> tok)
      out.push_back(tok);
    return out;
  }

  template<class T, class>
  struct Parser
  {
    static T parse(const std::string& str)
    {
      T val;
      std::istringstream s(str);
      s.imbue(std::locale::classic());
      s >> val;
      if (!s)
        DUNE_THROW(RangeError, " as a " << typeid(T).name());
      char dummy;
      s >> dummy;
      if (!s.eof() && s)
        DUNE_THROW(RangeError,
                   " as a " << typeid(T).name() << " (extra characters)");
      return val;
    }
  };

  template<class It>
  static void parseRange(const std::string& str, It begin, It end)
  {
    typedef typename std::iterator_traits<It>::value_type Value;
    const auto toks = split(str);
    if ((std::ptrdiff_t)toks.size() != std::distance(begin, end))
      DUNE_THROW(RangeError, "as a range of " << std::distance(begin, end)
                                              << " items (" << toks.size()
                                              << " items given)");
    std::size_t i = 0;
    for (It it = begin; it != end; ++it, ++i)
      *it = Parser<Value>::parse(toks[i]);
  }

  std::string _prefix;
  KeyVector _valueKeys, _subKeys;
  std::map<std::string, std::string> _values;
  std::map<std::string, ParameterTree> _subs;
};

template<>
struct ParameterTree::Parser<std::string, void>
{
  static std::string parse(const std::string& str) { return rtrim(ltrim(str)); }
};

template<>
struct ParameterTree::Parser<bool, void>
{
  static bool parse(const std::string& str)
  {
    std::string ret = rtrim(ltrim(str));
    std::transform(ret.begin(), ret.end(), ret.begin(), [](unsigned char c) {
      return (char)std::tolower(c);
    });
    if (ret == "yes" || ret == "true")
      return true;
    if (ret == "no" || ret == "false")
      return false;
    return Parser<int>::parse(ret) != 0;
  }
};

template<class T, int n>
struct ParameterTree::Parser<FieldVector<T, n>, void>
{
  static FieldVector<T, n> parse(const std::string& str)
  {
    FieldVector<T, n> val;
    parseRange(str, val.begin(), val.end());
    return val;
  }
};

template<class T, std::size_t n>
struct ParameterTree::Parser<std::array<T, n>, void>
{
  static std::array<T, n> parse(const std::string& str)
  {
    std::array<T, n> val;
    parseRange(str, val.begin(), val.end());
    return val;
  }
};

template<class T, class A>
struct ParameterTree::Parser<std::vector<T, A>, void>
{
  static std::vector<T, A> parse(const std::string& str)
  {
    std::vector<T, A> vec;
    for (const auto& tok : split(str))
      vec.push_back(Parser<T>::parse(tok));
    return vec;
  }
};

} // namespace Dune
#endif
