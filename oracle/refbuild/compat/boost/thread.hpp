// TEST INFRASTRUCTURE ONLY -- boost::thread / mutex / condition_variable / unique_lock mapped onto <thread> etc.
#ifndef DSO_COMPAT_BOOST_THREAD
#define DSO_COMPAT_BOOST_THREAD
#include <thread>
#include <mutex>
#include <condition_variable>
#include <functional>
#include "bind.hpp"
namespace boost
{
typedef std::mutex mutex;
typedef std::condition_variable condition_variable;
template <class M> using unique_lock = std::unique_lock<M>;
template <class M> using lock_guard = std::lock_guard<M>;
class thread
{
public:
	std::thread t;
	thread() {}
	template <class F, class... A> explicit thread(F f, A... a) : t(f, a...) {}
	thread(thread&& o) : t(std::move(o.t)) {}
	thread& operator=(thread&& o) { t = std::move(o.t); return *this; }
	void join() { if (t.joinable()) t.join(); }
	bool joinable() const { return t.joinable(); }
	void detach() { t.detach(); }
};
}
#endif
