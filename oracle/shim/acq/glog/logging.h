// Test-infrastructure stub (NOT product code): glog is not installed here.
#ifndef ORACLE_SHIM_ACQ_GLOG_H
#define ORACLE_SHIM_ACQ_GLOG_H
#include <iostream>
struct OracleNullLog
{
    template <typename T>
    OracleNullLog& operator<<(const T&) { return *this; }
    OracleNullLog& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define LOG(x) OracleNullLog()
#define DLOG(x) OracleNullLog()
#define VLOG(x) OracleNullLog()
#endif
