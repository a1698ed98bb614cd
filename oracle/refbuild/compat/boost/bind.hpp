// TEST INFRASTRUCTURE ONLY -- boost::bind / boost::function mapped onto <functional>; the global _1.._4
// placeholders of boost/bind.hpp come from std::placeholders.
#ifndef DSO_COMPAT_BOOST_BIND
#define DSO_COMPAT_BOOST_BIND
#include <functional>
using namespace std::placeholders;
namespace boost
{
template <class... A> auto bind(A&&... a) -> decltype(std::bind(std::forward<A>(a)...)) { return std::bind(std::forward<A>(a)...); }
template <class S> using function = std::function<S>;
}
#endif
