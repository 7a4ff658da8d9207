// TEST INFRASTRUCTURE ONLY.  Stand-in for the easylogging++ header the reference's benchmark includes
// (benchmark/gapped-matching/include/logging.hpp:10; the external/easyloggingpp submodule is not vendored under
// /root/reference): just enough for logging.hpp / utils.hpp to compile, every log statement a no-op.
#pragma once
#include <string>
#define INITIALIZE_EASYLOGGINGPP
#define START_EASYLOGGINGPP(argc, argv) ((void)(argc), (void)(argv))
namespace el {
enum class LoggingFlag { LogDetailedCrashReason, DisableApplicationAbortOnFatalLog, ColoredTerminalOutput, MultiLoggerSupport, HierarchicalLogging };
enum class ConfigurationType { Enabled, Format, ToFile, ToStandardOutput };
struct Configurations {
    void setGlobally(ConfigurationType, const std::string&) {}
};
struct Loggers {
    static void addFlag(LoggingFlag) {}
    static void reconfigureAllLoggers(const Configurations&) {}
};
struct NullLog {
    template <class T>
    NullLog& operator<<(const T&) { return *this; }
};
}  // namespace el
#define LOG(level) el::NullLog()
