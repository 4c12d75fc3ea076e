// Minimal stand-in for Dune::FieldVector (dune-common absent here).  Dense
// fixed-size vector; FieldVector<K,1> converts implicitly to and from K, which
// the random field code relies on (randomfield.hh:594-600 of the reference).
#ifndef PF_STANDIN_DUNE_COMMON_FVECTOR_HH
#define PF_STANDIN_DUNE_COMMON_FVECTOR_HH

#include <array>
#include <cmath>
#include <cstddef>
#include <istream>
#include <ostream>
#include <type_traits>

namespace Dune {

template<class K, int SIZE>
class FieldVector
{
public:
  using value_type = K;
  using field_type = K;
  using size_type = std::size_t;
  enum { dimension = SIZE };

  constexpr FieldVector() : _data{} {}
  constexpr FieldVector(const K& k) { for (auto& v : _data) v = k; }
  FieldVector(std::initializer_list<K> l)
  {
    std::size_t i = 0;
    for (const auto& v : l)
      if (i < SIZE) _data[i++] = v;
    for (; i < SIZE; i++) _data[i] = K();
  }
  template<class K2>
  FieldVector(const FieldVector<K2, SIZE>& o)
  {
    for (int i = 0; i < SIZE; i++) _data[i] = o[i];
  }

  static constexpr size_type size() { return SIZE; }
  K& operator[](size_type i) { return _data[i]; }
  const K& operator[](size_type i) const { return _data[i]; }
  K* data() { return _data.data(); }
  const K* data() const { return _data.data(); }
  auto begin() { return _data.begin(); }
  auto end() { return _data.end(); }
  auto begin() const { return _data.begin(); }
  auto end() const { return _data.end(); }

  FieldVector& operator=(const K& k)
  {
    for (auto& v : _data) v = k;
    return *this;
  }
  FieldVector& operator+=(const FieldVector& o)
  {
    for (int i = 0; i < SIZE; i++) _data[i] += o._data[i];
    return *this;
  }
  FieldVector& operator-=(const FieldVector& o)
  {
    for (int i = 0; i < SIZE; i++) _data[i] -= o._data[i];
    return *this;
  }
  FieldVector& operator*=(const K& k)
  {
    for (auto& v : _data) v *= k;
    return *this;
  }
  FieldVector& operator/=(const K& k)
  {
    for (auto& v : _data) v /= k;
    return *this;
  }
  FieldVector& axpy(const K& a, const FieldVector& x)
  {
    for (int i = 0; i < SIZE; i++) _data[i] += a * x._data[i];
    return *this;
  }
  K operator*(const FieldVector& o) const
  {
    K r = K();
    for (int i = 0; i < SIZE; i++) r += _data[i] * o._data[i];
    return r;
  }
  K dot(const FieldVector& o) const { return (*this) * o; }
  K two_norm2() const { return (*this) * (*this); }
  K two_norm() const { return std::sqrt(two_norm2()); }
  K one_norm() const
  {
    K r = K();
    for (const auto& v : _data) r += std::abs(v);
    return r;
  }
  K infinity_norm() const
  {
    K r = K();
    for (const auto& v : _data) r = std::abs(v) > r ? std::abs(v) : r;
    return r;
  }
  bool operator==(const FieldVector& o) const { return _data == o._data; }
  bool operator!=(const FieldVector& o) const { return !(*this == o); }

private:
  std::array<K, SIZE> _data;
};

// FieldVector<K,1> behaves like a scalar: implicit conversion both ways
template<class K>
class FieldVector<K, 1>
{
public:
  using value_type = K;
  using field_type = K;
  using size_type = std::size_t;
  enum { dimension = 1 };

  constexpr FieldVector() : _data() {}
  constexpr FieldVector(const K& k) : _data(k) {}
  template<class K2, class = std::enable_if_t<std::is_arithmetic<K2>::value>>
  constexpr FieldVector(const K2& k) : _data(k) {}
  template<class K2>
  FieldVector(const FieldVector<K2, 1>& o) : _data(o[0]) {}

  static constexpr size_type size() { return 1; }
  K& operator[](size_type) { return _data; }
  const K& operator[](size_type) const { return _data; }
  K* data() { return &_data; }
  const K* data() const { return &_data; }
  K* begin() { return &_data; }
  K* end() { return &_data + 1; }
  const K* begin() const { return &_data; }
  const K* end() const { return &_data + 1; }

  FieldVector& operator=(const K& k) { _data = k; return *this; }
  FieldVector& operator+=(const FieldVector& o) { _data += o._data; return *this; }
  FieldVector& operator-=(const FieldVector& o) { _data -= o._data; return *this; }
  FieldVector& operator+=(const K& k) { _data += k; return *this; }
  FieldVector& operator-=(const K& k) { _data -= k; return *this; }
  FieldVector& operator*=(const K& k) { _data *= k; return *this; }
  FieldVector& operator/=(const K& k) { _data /= k; return *this; }
  FieldVector& axpy(const K& a, const FieldVector& x)
  {
    _data += a * x._data;
    return *this;
  }
  K dot(const FieldVector& o) const { return _data * o._data; }
  K two_norm2() const { return _data * _data; }
  K two_norm() const { return std::abs(_data); }
  K one_norm() const { return std::abs(_data); }
  K infinity_norm() const { return std::abs(_data); }

  operator K&() { return _data; }
  operator const K&() const { return _data; }

private:
  K _data;
};

template<class K, int SIZE>
FieldVector<K, SIZE>
operator+(FieldVector<K, SIZE> a, const FieldVector<K, SIZE>& b)
{
  return a += b;
}
template<class K, int SIZE>
FieldVector<K, SIZE>
operator-(FieldVector<K, SIZE> a, const FieldVector<K, SIZE>& b)
{
  return a -= b;
}
// scalar-like mixed arithmetic for FieldVector<K,1>
#define PF_FV1_BINOP(op)                                                       \
  template<class K>                                                            \
  K operator op(const FieldVector<K, 1>& a, const K& b) { return a[0] op b; }  \
  template<class K>                                                            \
  K operator op(const K& a, const FieldVector<K, 1>& b) { return a op b[0]; }
PF_FV1_BINOP(+)
PF_FV1_BINOP(-)
PF_FV1_BINOP(*)
PF_FV1_BINOP(/)
#undef PF_FV1_BINOP
template<class K>
K operator*(const FieldVector<K, 1>& a, const FieldVector<K, 1>& b)
{
  return a[0] * b[0];
}
#define PF_FV1_CMP(op)                                                         \
  template<class K>                                                            \
  bool operator op(const FieldVector<K, 1>& a, const FieldVector<K, 1>& b)     \
  { return a[0] op b[0]; }                                                     \
  template<class K>                                                            \
  bool operator op(const FieldVector<K, 1>& a, const K& b) { return a[0] op b; } \
  template<class K>                                                            \
  bool operator op(const K& a, const FieldVector<K, 1>& b) { return a op b[0]; }
PF_FV1_CMP(==)
PF_FV1_CMP(!=)
PF_FV1_CMP(<)
PF_FV1_CMP(>)
PF_FV1_CMP(<=)
PF_FV1_CMP(>=)
#undef PF_FV1_CMP

template<class K, int SIZE>
std::ostream&
operator<<(std::ostream& s, const FieldVector<K, SIZE>& v)
{
  for (int i = 0; i < SIZE; i++) s << (i ? " " : "") << v[i];
  return s;
}
template<class K, int SIZE>
std::istream&
operator>>(std::istream& s, FieldVector<K, SIZE>& v)
{
  for (int i = 0; i < SIZE; i++) s >> v[i];
  return s;
}

} // namespace Dune
#endif
