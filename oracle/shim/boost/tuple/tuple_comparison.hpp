// TEST INFRASTRUCTURE ONLY -- std::tuple already has comparison operators.
#include <tuple>
