// SSPPException — the planner's error type (reference include/sspp/ssppexception.h:27-35): a
// std::exception carrying a message.  Non-zero C-ABI statuses are turned into this.
#pragma once
#include <exception>
#include <string>
class SSPPException : public std::exception {
 public:
  explicit SSPPException(const char* msg) : msg_(msg ? msg : "") {}
  explicit SSPPException(const std::string& msg) : msg_(msg) {}
  const char* what() const noexcept override { return msg_.c_str(); }
 private:
  std::string msg_;
};
