// Minimal stand-in for dune-common's exception classes (dune-common is not
// available in this environment).  Same names and
// semantics as <dune/common/exceptions.hh>: Dune::Exception derives from
// std::exception, DUNE_THROW(Type, stream-expression) throws with a message.
#ifndef PF_STANDIN_DUNE_COMMON_EXCEPTIONS_HH
#define PF_STANDIN_DUNE_COMMON_EXCEPTIONS_HH

#include <exception>
#include <sstream>
#include <string>

namespace Dune {

class Exception : public std::exception
{
public:
  Exception() = default;
  void message(const std::string& msg) { _message = msg; }
  const char* what() const noexcept override { return _message.c_str(); }

private:
  std::string _message;
};

inline std::ostream&
operator<<(std::ostream& stream, const Exception& e)
{
  return stream << e.what();
}

class IOError : public Exception {};
class MathError : public Exception {};
class RangeError : public Exception {};
class NotImplemented : public Exception {};
class InvalidStateException : public Exception {};
class ParameterTreeParserError : public RangeError {};

} // namespace Dune

#define DUNE_THROW(E, m)                                                       \
  do {                                                                         \
    E th__ex;                                                                  \
    std::ostringstream th__out;                                                \
    th__out << #E << " [" << __func__ << ":" << __FILE__ << ":" << __LINE__    \
            << "]: " << m;                                                     \
    th__ex.message(th__out.str());                                             \
    throw th__ex;                                                              \
  } while (0)

#endif
