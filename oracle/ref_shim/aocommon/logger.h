// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/logger.h (ours, nothing copied):
// stream-style log writers per level, a LogReceiver base class with a virtual Output(), and a
// global Logger whose output is discarded unless RDL_REF_VERBOSE is set in the environment.
#ifndef AOCOMMON_LOGGER_H_
#define AOCOMMON_LOGGER_H_

#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>

namespace aocommon {

class Logger {
 public:
  enum LoggerLevel { kNoLevel = 5, kFatalLevel = 4, kErrorLevel = 3, kWarningLevel = 2,
                     kInfoLevel = 1, kDebugLevel = 0 };
  enum VerbosityLevel { kQuietVerbosity, kNormalVerbosity, kVerboseVerbosity };

  template <LoggerLevel Level>
  class LogWriter {
   public:
    template <typename T>
    LogWriter& operator<<(const T& v) {
      if (Enabled()) std::cout << v;
      return *this;
    }
    LogWriter& operator<<(std::ostream& (*f)(std::ostream&)) {
      if (Enabled()) std::cout << f;
      return *this;
    }
    void Flush() {
      if (Enabled()) std::cout.flush();
    }
  };
  static bool Enabled() {
    static const bool on = std::getenv("RDL_REF_VERBOSE") != nullptr;
    return on;
  }
  static bool IsVerbose() { return false; }
  static void SetVerbosity(VerbosityLevel) {}
  static inline LogWriter<kDebugLevel> Debug;
  static inline LogWriter<kInfoLevel> Info;
  static inline LogWriter<kWarningLevel> Warn;
  static inline LogWriter<kErrorLevel> Error;
  static inline LogWriter<kFatalLevel> Fatal;
};

class LogReceiver {
 public:
  template <Logger::LoggerLevel Level>
  class LevelReceiver {
   public:
    explicit LevelReceiver(LogReceiver& parent) : parent_(&parent) {}
    LevelReceiver(const LevelReceiver&) = default;
    template <typename T>
    LevelReceiver& operator<<(const T& v) {
      std::ostringstream s;
      s << v;
      parent_->Output(Level, s.str());
      return *this;
    }
    LevelReceiver& operator<<(const std::string& s) {
      parent_->Output(Level, s);
      return *this;
    }
    LevelReceiver& operator<<(const char* s) {
      parent_->Output(Level, std::string(s));
      return *this;
    }
    LevelReceiver& operator<<(char c) {
      parent_->Output(Level, std::string(1, c));
      return *this;
    }
    void Flush() {}
    void Rebind(LogReceiver& parent) { parent_ = &parent; }

   private:
    LogReceiver* parent_;
  };

  LogReceiver() : Fatal(*this), Error(*this), Warn(*this), Info(*this), Debug(*this) {}
  LogReceiver(const LogReceiver&) : LogReceiver() {}
  LogReceiver(LogReceiver&&) noexcept : LogReceiver() {}
  LogReceiver& operator=(const LogReceiver&) { return *this; }
  LogReceiver& operator=(LogReceiver&&) noexcept { return *this; }
  virtual ~LogReceiver() = default;

  LevelReceiver<Logger::kFatalLevel> Fatal;
  LevelReceiver<Logger::kErrorLevel> Error;
  LevelReceiver<Logger::kWarningLevel> Warn;
  LevelReceiver<Logger::kInfoLevel> Info;
  LevelReceiver<Logger::kDebugLevel> Debug;

 protected:
  virtual void Output(enum Logger::LoggerLevel level, const std::string& str) = 0;
  void Forward(enum Logger::LoggerLevel /*level*/, const std::string& str) {
    if (Logger::Enabled()) std::cout << str;
  }
};

class ForwardingLogReceiver final : public LogReceiver {
 protected:
  void Output(enum Logger::LoggerLevel level, const std::string& str) override {
    Forward(level, str);
  }
};

}  // namespace aocommon
#endif
