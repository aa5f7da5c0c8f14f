// Minimal stand-in for <boost/thread/recursive_mutex.hpp> (oracle build only).
#pragma once
#include <mutex>
namespace boost {
class recursive_mutex : public std::recursive_mutex {
 public:
  class scoped_lock {
   public:
    explicit scoped_lock(recursive_mutex &m) : m_(m) { m_.lock(); }
    ~scoped_lock() { m_.unlock(); }
   private:
    recursive_mutex &m_;
  };
};
}  // namespace boost
