// oracle/shim/base/mutex.h -- TEST INFRASTRUCTURE ONLY.
// Stand-in for glog's internal <base/mutex.h>, which two reference files
// include (integrators/photonbeam.cpp:43, integrators/photonmap.cpp:37).
#ifndef BRE_ORACLE_SHIM_BASE_MUTEX_H
#define BRE_ORACLE_SHIM_BASE_MUTEX_H
#include <mutex>
class Mutex {
  public:
    void Lock() { m.lock(); }
    void Unlock() { m.unlock(); }
  private:
    std::mutex m;
};
class MutexLock {
  public:
    explicit MutexLock(Mutex *mu) : mu(mu) { mu->Lock(); }
    ~MutexLock() { mu->Unlock(); }
  private:
    Mutex *mu;
};
#endif
