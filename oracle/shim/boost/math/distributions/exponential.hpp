/* TEST INFRASTRUCTURE ONLY: boost::math::exponential_distribution and its quantile (-log1p(-p) / lambda), which is all the reference's
 * adapters use (calculate_threshold). */
#ifndef GB_ORACLE_SHIM_BOOST_MATH_EXPONENTIAL_HPP
#define GB_ORACLE_SHIM_BOOST_MATH_EXPONENTIAL_HPP
#include <cmath>
namespace boost
{
namespace math
{
template <class T = double>
class exponential_distribution
{
public:
    explicit exponential_distribution(T lambda = 1) : lambda_(lambda) {}
    T lambda() const { return lambda_; }

private:
    T lambda_;
};
template <class T>
inline T quantile(const exponential_distribution<T>& d, const T& p)
{
    return -std::log1p(-p) / d.lambda();
}
}  // namespace math
}  // namespace boost
#endif
