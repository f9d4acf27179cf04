// TEST INFRASTRUCTURE ONLY -- third/mesh/*.cpp include boost::tuple for one unused typedef.
#include <tuple>
namespace boost { template <typename... A> using tuple = std::tuple<A...>; }
