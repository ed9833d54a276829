// oracle/shim: TEST INFRASTRUCTURE ONLY. An in-process stand-in for the roscpp client library, so that the reference's
// node/*.cpp compile UNMODIFIED into oracle/_ref and can be driven by the harness: parameters come from a table, topics
// deliver synchronously to the registered member callbacks, published messages are kept (last one per topic) for the
// harness to read, services and timers are callable by name / index, time is a settable clock. Nothing here computes
// anything the parity tests compare; it only routes the reference's own code.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <ostream>
#include <string>
#include <utility>
#include <vector>
namespace ros
{
namespace shim
{
typedef std::function<void(const std::shared_ptr<const void>&)> AnyCallback;
struct Bus
{
  std::map<std::string, double> num_params;
  std::map<std::string, std::string> str_params;
  struct Sub
  {
    long id;
    AnyCallback cb;
  };
  std::map<std::string, std::vector<Sub>> subs;
  std::map<std::string, std::shared_ptr<const void>> last_msg;
  std::map<std::string, long> msg_count;
  std::map<std::string, std::function<bool()>> services;
  struct TimerRec
  {
    long id;
    double period;
    std::function<void()> fire;
  };
  std::vector<TimerRec> timers;
  std::function<void()> reconfigure;  // set by dynamic_reconfigure::Server::setCallback
  double now = 0.0;
  long next_id = 1;

  long subscribe(const std::string& topic, AnyCallback cb)
  {
    subs[topic].push_back({next_id, std::move(cb)});
    return next_id++;
  }
  void unsubscribe(const std::string& topic, long id)
  {
    std::vector<Sub>& v = subs[topic];
    for (size_t k = 0; k < v.size(); k++)
      if (v[k].id == id)
      {
        v.erase(v.begin() + k);
        return;
      }
  }
  int deliver(const std::string& topic, const std::shared_ptr<const void>& msg)
  {
    std::vector<Sub> v = subs[topic];  // a callback may (un)subscribe
    for (Sub& s : v)
      s.cb(msg);
    return static_cast<int>(v.size());
  }
  void remove_timer(long id)
  {
    for (size_t k = 0; k < timers.size(); k++)
      if (timers[k].id == id)
      {
        timers.erase(timers.begin() + k);
        return;
      }
  }
  void reset()
  {
    num_params.clear();
    str_params.clear();
    last_msg.clear();
    msg_count.clear();
    now = 0.0;
  }
};
inline Bus& bus()
{
  static Bus b;
  return b;
}
}  // namespace shim

class Duration
{
public:
  Duration() : s_(0.0) {}
  explicit Duration(double t) : s_(t) {}
  Duration& fromSec(double t)
  {
    s_ = t;
    return *this;
  }
  double toSec() const { return s_; }
  bool operator>(const Duration& o) const { return s_ > o.s_; }
  bool operator<(const Duration& o) const { return s_ < o.s_; }
private:
  double s_;
};
inline std::ostream& operator<<(std::ostream& os, const Duration& d) { return os << d.toSec(); }

class Time
{
public:
  uint32_t sec, nsec;
  Time() : sec(0), nsec(0) {}
  explicit Time(double t) { fromSec(t); }
  Time& fromSec(double t)
  {
    sec = static_cast<uint32_t>(t);
    nsec = static_cast<uint32_t>((t - sec) * 1e9 + 0.5);
    return *this;
  }
  double toSec() const { return static_cast<double>(sec) + 1e-9 * static_cast<double>(nsec); }
  static Time now() { return Time(shim::bus().now); }
  Time operator+(const Duration& d) const { return Time(toSec() + d.toSec()); }
  Duration operator-(const Time& o) const { return Duration(toSec() - o.toSec()); }
  bool operator==(const Time& o) const { return sec == o.sec && nsec == o.nsec; }
};

struct TimerEvent
{
};
class CallbackQueue
{
};
class AsyncSpinner
{
public:
  AsyncSpinner(uint32_t, CallbackQueue* = nullptr) {}
  void start() {}
  void stop() {}
};

class Publisher
{
public:
  Publisher() {}
  explicit Publisher(const std::string& topic) : topic_(topic) {}
  template <typename M>
  void publish(const M& msg) const
  {
    if (topic_.empty())
      return;
    shim::Bus& b = shim::bus();
    std::shared_ptr<const M> copy = std::make_shared<const M>(msg);
    b.last_msg[topic_] = copy;
    b.msg_count[topic_]++;
    b.deliver(topic_, copy);
  }
  std::string getTopic() const { return topic_; }
private:
  std::string topic_;
};
class SingleSubscriberPublisher
{
public:
  template <typename M>
  void publish(const M&) const {}
};

// handles share an implementation; the last copy going away unsubscribes / removes, as in roscpp
class Subscriber
{
public:
  Subscriber() {}
  Subscriber(const std::string& topic, long id) : impl_(new Impl{topic, id}, &Subscriber::drop) {}
  std::string getTopic() const { return impl_ ? impl_->topic : std::string(); }
private:
  struct Impl
  {
    std::string topic;
    long id;
  };
  static void drop(Impl* p)
  {
    shim::bus().unsubscribe(p->topic, p->id);
    delete p;
  }
  std::shared_ptr<Impl> impl_;
};
class ServiceServer
{
public:
  ServiceServer() {}
  explicit ServiceServer(const std::string& name) : impl_(new std::string(name), &ServiceServer::drop) {}
private:
  static void drop(std::string* p)
  {
    shim::bus().services.erase(*p);
    delete p;
  }
  std::shared_ptr<std::string> impl_;
};
class Timer
{
public:
  Timer() {}
  explicit Timer(long id) : impl_(new long(id), &Timer::drop) {}
  void setPeriod(const Duration& d)
  {
    if (!impl_)
      return;
    for (shim::Bus::TimerRec& t : shim::bus().timers)
      if (t.id == *impl_)
        t.period = d.toSec();
  }
  void start() {}
  void stop() {}
private:
  static void drop(long* p)
  {
    shim::bus().remove_timer(*p);
    delete p;
  }
  std::shared_ptr<long> impl_;
};

class NodeHandle
{
public:
  NodeHandle() {}
  explicit NodeHandle(const std::string&) {}  // every parameter the node reads is private ("~"); one flat table
  void setCallbackQueue(CallbackQueue*) {}

  bool param(const std::string& name, int& val, const int& def) const { return num(name, val, def); }
  bool param(const std::string& name, double& val, const double& def) const { return num(name, val, def); }
  bool param(const std::string& name, bool& val, const bool& def) const
  {
    double v;
    const bool found = num(name, v, def ? 1.0 : 0.0);
    val = v != 0.0;
    return found;
  }
  bool param(const std::string& name, std::string& val, const std::string& def) const
  {
    const std::map<std::string, std::string>& t = shim::bus().str_params;
    std::map<std::string, std::string>::const_iterator it = t.find(name);
    val = it == t.end() ? def : it->second;
    return it != t.end();
  }

  template <class M, class T>
  Subscriber subscribe(const std::string& topic, uint32_t, void (T::*fp)(const std::shared_ptr<M const>&), T* obj)
  {
    long id = shim::bus().subscribe(topic, [fp, obj](const std::shared_ptr<const void>& m) {
      (obj->*fp)(std::static_pointer_cast<const M>(m));
    });
    return Subscriber(topic, id);
  }
  template <class M>
  Publisher advertise(const std::string& topic, uint32_t, bool = false)
  {
    return Publisher(topic);
  }
  template <class T, class Req, class Res>
  ServiceServer advertiseService(const std::string& name, bool (T::*fp)(Req&, Res&), T* obj)
  {
    shim::bus().services[name] = [fp, obj]() {
      Req req;
      Res res;
      return (obj->*fp)(req, res);
    };
    return ServiceServer(name);
  }
  Timer createTimer(Duration period, const std::function<void(const TimerEvent&)>& cb)
  {
    shim::Bus& b = shim::bus();
    const long id = b.next_id++;
    b.timers.push_back({id, period.toSec(), [cb]() { cb(TimerEvent()); }});
    return Timer(id);
  }
private:
  template <typename T>
  static bool num(const std::string& name, T& val, double def)
  {
    const std::map<std::string, double>& t = shim::bus().num_params;
    std::map<std::string, double>::const_iterator it = t.find(name);
    val = static_cast<T>(it == t.end() ? def : it->second);
    return it != t.end();
  }
};
}  // namespace ros
