// TEST INFRASTRUCTURE ONLY -- boost::bind / boost::function mapped onto <functional>; the global _1.._4
// placeholders of boost/bind.hpp come from std::placeholders.
#ifndef DSO_COMPAT_BOOST_FUNCTION
#define DSO_COMPAT_BOOST_FUNCTION
#include <functional>
#include "bind.hpp"

namespace boost
{

template <class S> using function = std::function<S>;
}
#endif
