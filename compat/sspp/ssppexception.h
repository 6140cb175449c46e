// Stand-in for sspp/include/sspp/ssppexception.h.
#ifndef EXCEPTIONHANDLER_H_
#define EXCEPTIONHANDLER_H_
#include <exception>
class SSPPException : public std::exception {
  const char* m_;
 public:
  explicit SSPPException(char* m) : m_(m) {}
  virtual const char* what() const throw() { return m_; }
};
#endif
