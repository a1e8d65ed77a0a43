// Test-infrastructure stub (NOT product code): gr::thread::mutex / scoped_lock are boost::mutex / boost::mutex::scoped_lock in GNU Radio
#ifndef ORACLE_SHIM_ACQ_GR_THREAD_H
#define ORACLE_SHIM_ACQ_GR_THREAD_H
#include <mutex>
namespace gr
{
namespace thread
{
typedef std::recursive_mutex mutex;
typedef std::unique_lock<std::recursive_mutex> scoped_lock;
}  // namespace thread
}  // namespace gr
#endif
