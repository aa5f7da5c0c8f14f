// Minimal stand-in for <boost/threadpool.hpp> (oracle build only): a pool of
// zero threads that runs scheduled work inline. The reference only uses a pool
// when island_threads / quickstep threads > 0, which the oracle never sets.
#pragma once
#include <cstddef>
#include <functional>
#include "boost/bind.hpp"
#include "boost/thread/mutex.hpp"
#include "boost/thread/recursive_mutex.hpp"
namespace boost { namespace threadpool {
class pool {
 public:
  explicit pool(std::size_t = 0) {}
  std::size_t size() const { return 0; }
  template <class F> bool schedule(F f) { f(); return true; }
  void wait() {}
};
}}  // namespace boost::threadpool
