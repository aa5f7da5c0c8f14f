// Minimal stand-in for <boost/shared_array.hpp> (oracle build only).
#pragma once
#include <cstddef>
#include <memory>
namespace boost {
template <class T> class shared_array {
 public:
  shared_array() {}
  explicit shared_array(T *p) : p_(p, std::default_delete<T[]>()) {}
  void reset(T *p = nullptr) { p_.reset(p, std::default_delete<T[]>()); }
  T &operator[](std::ptrdiff_t i) const { return p_.get()[i]; }
  T *get() const { return p_.get(); }
  explicit operator bool() const { return static_cast<bool>(p_); }
 private:
  std::shared_ptr<T> p_;
};
}  // namespace boost
