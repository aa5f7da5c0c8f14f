// Minimal stand-in for <boost/thread/mutex.hpp> so the reference's opende
// sources compile without Boost (test infrastructure only; see oracle/README.md).
#pragma once
#include <mutex>
namespace boost {
class mutex : public std::mutex {
 public:
  class scoped_lock {
   public:
    explicit scoped_lock(mutex &m) : m_(m) { m_.lock(); }
    ~scoped_lock() { m_.unlock(); }
   private:
    mutex &m_;
  };
};
}  // namespace boost
