// Parameter-server transport: a small TCP request/response protocol with a named variable store,
// lock-free (Hogwild) optimizer apply, named blocking queues and a ps-side checkpoint save.
//
// This is the native replacement for what the reference gets from tf.train.Server (in-process gRPC
// MasterService/WorkerService, reference trainer/task.py:118-122), the Send/Recv rendezvous behind
// every sess.run (trainer/task.py:236), the Apply* kernels with use_locking=False
// (trainer/task.py:138) and the shared_name FIFOQueue used by auto_stop_ps
// (auto_stop_ps/task.py:127-150, 270-272).
//
// Wire format (little endian):  request  = u32 len | u8 op | payload
//                               response = u32 len | i32 status | payload
#include <arpa/inet.h>
#include <errno.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <sys/uio.h>
#include <sys/time.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <thread>

#include "../../include/dtf_native.h"
#include "util.h"

namespace dtfn {

enum Op : uint8_t {
  OP_PING = 1, OP_INIT_VAR, OP_PULL, OP_LIST, OP_VAR_INFO, OP_PUSH_APPLY, OP_PUSH_PULL, OP_ASSIGN_ADD,
  OP_QUEUE_CREATE, OP_ENQUEUE, OP_DEQUEUE, OP_SAVE, OP_SHUTDOWN
};
enum { DT_FLOAT = 1, DT_INT32 = 3, DT_INT64 = 9 };

// ---------------------------------------------------------------- (de)serialisation
struct Writer {
  std::string s;
  void u8(uint8_t v) { s.push_back(static_cast<char>(v)); }
  void i32(int32_t v) { s.append(reinterpret_cast<const char*>(&v), 4); }
  void i64(int64_t v) { s.append(reinterpret_cast<const char*>(&v), 8); }
  void f32(float v) { s.append(reinterpret_cast<const char*>(&v), 4); }
  void str(const std::string& v) {
    i32(static_cast<int32_t>(v.size()));
    s.append(v);
  }
  // length, pad count, pad, payload: the payload starts at a multiple of 4 bytes from the start of the
  // message body, and both ends keep the body 4-byte aligned in memory, so float payloads can be
  // consumed in place
  void bytes(const void* p, size_t n) {
    i64(static_cast<int64_t>(n));
    const size_t pad = (4 - ((s.size() + 1) & 3)) & 3;
    u8(static_cast<uint8_t>(pad));
    s.append(pad, '\0');
    s.append(static_cast<const char*>(p), n);
  }
};
struct Reader {
  const char* p;
  const char* end;
  bool ok = true;
  Reader(const char* b, size_t n) : p(b), end(b + n) {}
  bool need(size_t n) {
    if (static_cast<size_t>(end - p) < n) ok = false;
    return ok;
  }
  uint8_t u8() { return need(1) ? static_cast<uint8_t>(*p++) : 0; }
  int32_t i32() {
    int32_t v = 0;
    if (need(4)) { memcpy(&v, p, 4); p += 4; }
    return v;
  }
  int64_t i64() {
    int64_t v = 0;
    if (need(8)) { memcpy(&v, p, 8); p += 8; }
    return v;
  }
  float f32() {
    float v = 0;
    if (need(4)) { memcpy(&v, p, 4); p += 4; }
    return v;
  }
  std::string str() {
    const int32_t n = i32();
    if (n < 0 || !need(static_cast<size_t>(n))) { ok = false; return std::string(); }
    std::string v(p, n);
    p += n;
    return v;
  }
  const char* bytes(size_t* n) {
    const int64_t len = i64();
    const size_t pad = u8();
    if (pad > 3 || !need(pad)) { ok = false; *n = 0; return nullptr; }
    p += pad;
    if (len < 0 || !need(static_cast<size_t>(len))) { ok = false; *n = 0; return nullptr; }
    const char* q = p;
    p += len;
    *n = static_cast<size_t>(len);
    return q;
  }
};

static bool send_all(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n) {
    const ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
    if (w <= 0) {
      if (w < 0 && errno == EINTR) continue;
      return false;
    }
    p += w;
    n -= static_cast<size_t>(w);
  }
  return true;
}
static bool recv_all(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n) {
    const ssize_t r = ::recv(fd, p, n, 0);
    if (r <= 0) {
      if (r < 0 && errno == EINTR) continue;
      return false;
    }
    p += r;
    n -= static_cast<size_t>(r);
  }
  return true;
}

// header + body with one gather write: the body (up to hundreds of MB of gradients or values) is never
// copied into a contiguous frame
static bool send_frame(int fd, const void* hdr, size_t hn, const std::string& body) {
  struct iovec iov[2];
  iov[0].iov_base = const_cast<void*>(hdr);
  iov[0].iov_len = hn;
  iov[1].iov_base = const_cast<char*>(body.data());
  iov[1].iov_len = body.size();
  int first = 0;
  while (first < 2) {
    if (iov[first].iov_len == 0) {
      ++first;
      continue;
    }
    struct msghdr m{};
    m.msg_iov = iov + first;
    m.msg_iovlen = static_cast<size_t>(2 - first);
    const ssize_t w = ::sendmsg(fd, &m, MSG_NOSIGNAL);
    if (w <= 0) {
      if (w < 0 && errno == EINTR) continue;
      return false;
    }
    size_t left = static_cast<size_t>(w);
    while (left > 0 && first < 2) {
      const size_t take = left < iov[first].iov_len ? left : iov[first].iov_len;
      iov[first].iov_base = static_cast<char*>(iov[first].iov_base) + take;
      iov[first].iov_len -= take;
      left -= take;
      if (iov[first].iov_len == 0) ++first;
    }
  }
  return true;
}
static void tune_socket(int fd) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  int sz = 4 << 20;   // dense gradient pushes are megabytes: let a whole message sit in the socket buffers
  setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &sz, sizeof(sz));
  setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &sz, sizeof(sz));
}

// ---------------------------------------------------------------- store
struct Var {
  int dtype = DT_FLOAT;
  std::vector<long long> dims;
  std::vector<char> data;
  size_t count() const { return data.size() / (dtype == DT_INT64 ? 8 : 4); }
  float* f() { return reinterpret_cast<float*>(data.data()); }
};

struct Queue {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<int> items;
  int capacity = 0;
};

struct Store {
  std::shared_mutex mu;  // guards the maps only; values are updated without locks (Hogwild)
  std::map<std::string, std::unique_ptr<Var>> vars;
  std::mutex qmu;
  std::map<std::string, std::unique_ptr<Queue>> queues;

  Var* find(const std::string& name) {
    std::shared_lock<std::shared_mutex> l(mu);
    auto it = vars.find(name);
    return it == vars.end() ? nullptr : it->second.get();
  }
  // creates a float slot shaped like `like`, filled with `init`, if absent
  Var* slot(const std::string& name, const Var* like, float init) {
    if (Var* v = find(name)) return v;
    std::unique_lock<std::shared_mutex> l(mu);
    auto& p = vars[name];
    if (!p) {
      p = std::make_unique<Var>();
      p->dtype = DT_FLOAT;
      if (like) {
        p->dims = like->dims;
        p->data.resize(like->data.size());
      } else {
        p->data.resize(4);
      }
      float* f = p->f();
      for (size_t i = 0; i < p->count(); ++i) f[i] = init;
    }
    return p.get();
  }
  Queue* queue(const std::string& name, int capacity, bool create) {
    std::lock_guard<std::mutex> l(qmu);
    auto it = queues.find(name);
    if (it != queues.end()) return it->second.get();
    if (!create) return nullptr;
    auto q = std::make_unique<Queue>();
    q->capacity = capacity;
    Queue* r = q.get();
    queues[name] = std::move(q);
    return r;
  }
};

// TF-1 default hyper-parameters (the reference passes only learning_rate, trainer/task.py:42-53)
static const float kAdamB1 = 0.9f, kAdamB2 = 0.999f, kAdamEps = 1e-8f;
static const float kAdadeltaRho = 0.95f, kAdadeltaEps = 1e-8f;
static const float kRmsDecay = 0.9f, kRmsMomentum = 0.0f, kRmsEps = 1e-10f;
static const float kFtrlPower = -0.5f, kFtrlL1 = 0.f, kFtrlL2 = 0.f;

// The three places where variable storage is touched without a lock, by design (Hogwild, TF's
// use_locking=False).  They are separate non-inlined functions so that csrc/native/tsan.supp can name
// exactly these and ThreadSanitizer keeps checking everything else.
__attribute__((noinline)) static void copy_value_unlocked(Writer& w, const Var* v) { w.bytes(v->data.data(), v->data.size()); }
__attribute__((noinline)) static std::vector<char> snapshot_value_unlocked(const Var* v) { return v->data; }
__attribute__((noinline)) static void hogwild_read_beta_powers(Store* st, float* b1p, float* b2p) {
  *b1p = st->slot("beta1_power", nullptr, kAdamB1)->f()[0];
  *b2p = st->slot("beta2_power", nullptr, kAdamB2)->f()[0];
}
__attribute__((noinline)) static void hogwild_advance_beta_powers(Store* st) {
  st->find("beta1_power")->f()[0] *= kAdamB1;
  st->find("beta2_power")->f()[0] *= kAdamB2;
}

// float view of a gradient inside the receive buffer: used in place when it happens to be 4-byte
// aligned, copied into `tmp` otherwise
static const float* grad_view(const char* g, size_t nbytes, std::vector<float>& tmp) {
  if ((reinterpret_cast<uintptr_t>(g) & 3u) == 0) return reinterpret_cast<const float*>(g);
  tmp.resize(nbytes / 4);
  memcpy(tmp.data(), g, tmp.size() * 4);
  return tmp.data();
}

static int apply_one(Store* st, const std::string& name, int opt, float lr, const float* g, size_t n,
                     float b1p, float b2p) {
  Var* v = st->find(name);
  if (!v) {
    set_error("ps: variable '%s' does not exist", name.c_str());
    return -10;
  }
  if (v->dtype != DT_FLOAT || v->count() != n) {
    set_error("ps: gradient for '%s' has %zu elements, variable has %zu", name.c_str(), n, v->count());
    return -11;
  }
  switch (opt) {
    case DTFN_OPT_SGD: dtfn_apply_sgd(v->f(), g, n, lr); break;
    case DTFN_OPT_ADAGRAD: dtfn_apply_adagrad(v->f(), st->slot(name + "/Adagrad", v, 0.1f)->f(), g, n, lr); break;
    case DTFN_OPT_ADAM:
      dtfn_apply_adam(v->f(), st->slot(name + "/Adam", v, 0.f)->f(), st->slot(name + "/Adam_1", v, 0.f)->f(), g, n, lr,
                      kAdamB1, kAdamB2, kAdamEps, b1p, b2p);
      break;
    case DTFN_OPT_ADADELTA:
      dtfn_apply_adadelta(v->f(), st->slot(name + "/Adadelta", v, 0.f)->f(), st->slot(name + "/Adadelta_1", v, 0.f)->f(),
                          g, n, lr, kAdadeltaRho, kAdadeltaEps);
      break;
    case DTFN_OPT_FTRL:
      dtfn_apply_ftrl(v->f(), st->slot(name + "/Ftrl", v, 0.1f)->f(), st->slot(name + "/Ftrl_1", v, 0.f)->f(), g, n, lr,
                      kFtrlL1, kFtrlL2, kFtrlPower);
      break;
    case DTFN_OPT_RMSPROP:
      dtfn_apply_rmsprop(v->f(), st->slot(name + "/RMSProp", v, 1.0f)->f(), st->slot(name + "/RMSProp_1", v, 0.f)->f(), g,
                         n, lr, kRmsDecay, kRmsMomentum, kRmsEps);
      break;
    default:
      set_error("ps: unknown optimizer id %d", opt);
      return -12;
  }
  return 0;
}

}  // namespace dtfn

using namespace dtfn;

struct dtfn_server {
  int listen_fd = -1;
  int port = 0;
  std::atomic<bool> stopping{false};
  std::atomic<int> drop{0};
  Store store;
  std::thread accept_thread;
  // live connections by id (fd numbers are reused by the process, ids are not); a handler that ends
  // moves its id to `finished` so that the accept loop can join and forget its thread
  struct Conn {
    int fd = -1;
    std::thread th;
  };
  std::mutex conn_mu;
  std::map<long long, Conn> conns;
  std::vector<long long> finished;
  long long next_conn_id = 1;
  std::mutex join_mu;
  std::condition_variable join_cv;
  bool stopped = false;

  void handle(long long id, int fd);
  void reap_finished();
  int dispatch(uint8_t op, Reader& r, Writer& w);
  void request_stop() {
    {
      std::lock_guard<std::mutex> l(join_mu);
      stopped = true;
    }
    stopping = true;
    join_cv.notify_all();
  }
};

int dtfn_server::dispatch(uint8_t op, Reader& r, Writer& w) {
  switch (op) {
    case OP_PING: return 0;
    case OP_INIT_VAR: {
      const std::string name = r.str();
      const int dtype = r.i32();
      const int nd = r.i32();
      std::vector<long long> dims;
      for (int i = 0; i < nd && r.ok; ++i) dims.push_back(r.i64());
      const int overwrite = r.u8();
      size_t n;
      const char* data = r.bytes(&n);
      if (!r.ok) return -1;
      // dtype and payload size come from the wire: a variable whose bytes do not match its declared
      // shape would later be indexed past its end by the applies and the atomic step counter
      if (dtype != DT_FLOAT && dtype != DT_INT32 && dtype != DT_INT64) return -11;
      if (nd < 0 || nd > 8) return -1;
      unsigned long long count = 1;
      for (long long d : dims) {
        if (d < 0 || (d > 0 && count > (1ull << 40) / static_cast<unsigned long long>(d))) return -1;
        count *= static_cast<unsigned long long>(d);
      }
      if (count * (dtype == DT_INT64 ? 8u : 4u) != n) return -12;
      std::unique_lock<std::shared_mutex> l(store.mu);
      auto& p = store.vars[name];
      if (p && !overwrite) return 1;
      if (!p) p = std::make_unique<Var>();
      p->dtype = dtype;
      p->dims = dims;
      p->data.assign(data, data + n);
      return 0;
    }
    case OP_PULL: {
      const std::string name = r.str();
      if (!r.ok) return -1;
      Var* v = store.find(name);
      if (!v) return -10;
      copy_value_unlocked(w, v);
      return 0;
    }
    case OP_LIST: {
      std::string all;
      std::shared_lock<std::shared_mutex> l(store.mu);
      for (auto& kv : store.vars) {
        if (!all.empty()) all.push_back('\n');
        all += kv.first;
      }
      w.str(all);
      return 0;
    }
    case OP_VAR_INFO: {
      const std::string name = r.str();
      if (!r.ok) return -1;
      Var* v = store.find(name);
      if (!v) return -10;
      w.i32(v->dtype);
      w.i32(static_cast<int32_t>(v->dims.size()));
      for (long long d : v->dims) w.i64(d);
      w.i64(static_cast<int64_t>(v->data.size()));
      return 0;
    }
    case OP_PUSH_APPLY: {
      const std::string name = r.str();
      const int opt = r.i32();
      const float lr = r.f32();
      size_t n;
      const char* g = r.bytes(&n);
      if (!r.ok) return -1;
      float b1p = kAdamB1, b2p = kAdamB2;
      if (opt == DTFN_OPT_ADAM) {
        hogwild_read_beta_powers(&store, &b1p, &b2p);
      }
      std::vector<float> tmp;
      const int rc = apply_one(&store, name, opt, lr, grad_view(g, n, tmp), n / 4, b1p, b2p);
      if (rc == 0 && opt == DTFN_OPT_ADAM) {
        hogwild_advance_beta_powers(&store);
      }
      return rc;
    }
    case OP_PUSH_PULL: {
      const int k = r.i32();
      const int opt = r.i32();
      const float lr = r.f32();
      float b1p = kAdamB1, b2p = kAdamB2;
      if (opt == DTFN_OPT_ADAM && k > 0) {
        hogwild_read_beta_powers(&store, &b1p, &b2p);
      }
      std::vector<std::string> names;
      for (int i = 0; i < k && r.ok; ++i) {
        names.push_back(r.str());
        size_t n;
        const char* g = r.bytes(&n);
        if (!r.ok) return -1;
        std::vector<float> tmp;
        const int rc = apply_one(&store, names.back(), opt, lr, grad_view(g, n, tmp), n / 4, b1p, b2p);
        if (rc != 0) return rc;
      }
      if (opt == DTFN_OPT_ADAM && k > 0) {
        hogwild_advance_beta_powers(&store);
      }
      const std::string step_name = r.str();
      const int64_t inc = r.i64();
      const int want = r.u8();
      const int n_extra = r.i32();
      std::vector<std::string> extra;
      for (int i = 0; i < n_extra && r.ok; ++i) extra.push_back(r.str());
      if (!r.ok) return -1;
      int64_t step = -1;
      if (!step_name.empty()) {
        Var* sv = store.find(step_name);
        if (!sv) return -10;
        if (sv->dtype == DT_INT64 && sv->data.size() >= 8) {
          auto* a = reinterpret_cast<std::atomic<int64_t>*>(sv->data.data());
          step = a->fetch_add(inc) + inc;
        } else if (sv->dtype == DT_INT32 && sv->data.size() >= 4) {
          auto* a = reinterpret_cast<std::atomic<int32_t>*>(sv->data.data());
          step = a->fetch_add(static_cast<int32_t>(inc)) + static_cast<int32_t>(inc);
        } else {
          return -11;   // the step variable must be an integer scalar
        }
      }
      w.i64(step);
      if (want) {
        for (const std::string& nm : names) {
          Var* v = store.find(nm);
          copy_value_unlocked(w, v);
        }
        for (const std::string& nm : extra) {
          Var* v = store.find(nm);
          if (!v) return -10;
          copy_value_unlocked(w, v);
        }
      }
      return 0;
    }
    case OP_ASSIGN_ADD: {
      const std::string name = r.str();
      const int64_t delta = r.i64();
      if (!r.ok) return -1;
      Var* sv = store.find(name);
      if (!sv) return -10;
      int64_t out;
      if (sv->dtype == DT_INT64 && sv->data.size() >= 8) {
        auto* a = reinterpret_cast<std::atomic<int64_t>*>(sv->data.data());
        out = a->fetch_add(delta) + delta;
      } else if (sv->dtype == DT_INT32 && sv->data.size() >= 4) {
        auto* a = reinterpret_cast<std::atomic<int32_t>*>(sv->data.data());
        out = a->fetch_add(static_cast<int32_t>(delta)) + static_cast<int32_t>(delta);
      } else {
        return -11;
      }
      w.i64(out);
      return 0;
    }
    case OP_QUEUE_CREATE: {
      const std::string name = r.str();
      const int cap = r.i32();
      if (!r.ok) return -1;
      store.queue(name, cap, true);
      return 0;
    }
    case OP_ENQUEUE: {
      const std::string name = r.str();
      const int val = r.i32();
      if (!r.ok) return -1;
      // optional trailing field: how long the server may hold the request while the queue is full; the
      // client keeps it below its socket timeout, so no late reply can get out of step with a later request
      int wait_ms = 500;
      if (static_cast<size_t>(r.end - r.p) >= 4) wait_ms = r.i32();
      Queue* q = store.queue(name, 1 << 20, true);
      std::unique_lock<std::mutex> l(q->mu);
      const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(wait_ms);
      while (q->capacity > 0 && static_cast<int>(q->items.size()) >= q->capacity) {
        if (stopping) return -2;
        if (q->cv.wait_until(l, deadline) == std::cv_status::timeout &&
            static_cast<int>(q->items.size()) >= q->capacity)
          return -2;   // still full: the client retries
      }
      q->items.push_back(val);
      q->cv.notify_all();
      return 0;
    }
    case OP_DEQUEUE: {
      const std::string name = r.str();
      const int timeout_ms = r.i32();
      if (!r.ok) return -1;
      Queue* q = store.queue(name, 1 << 20, true);
      std::unique_lock<std::mutex> l(q->mu);
      const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
      while (q->items.empty()) {
        if (stopping) return -2;
        if (q->cv.wait_until(l, deadline) == std::cv_status::timeout && q->items.empty()) return -2;
      }
      w.i32(q->items.front());
      q->items.pop_front();
      q->cv.notify_all();
      return 0;
    }
    case OP_SAVE: {
      const std::string prefix = r.str();
      const int n_excl = r.i32();
      std::vector<std::string> excl;
      for (int i = 0; i < n_excl && r.ok; ++i) excl.push_back(r.str());
      if (!r.ok) return -1;
      dtfn_bundle_writer* bw = dtfn_bundle_writer_new(prefix.c_str());
      if (!bw) return -20;
      std::shared_lock<std::shared_mutex> l(store.mu);
      for (auto& kv : store.vars) {
        if (std::find(excl.begin(), excl.end(), kv.first) != excl.end()) continue;
        // snapshot without stopping writers: like TF, a checkpoint taken during Hogwild training is
        // not an atomic cut across variables
        std::vector<char> copy = snapshot_value_unlocked(kv.second.get());
        if (dtfn_bundle_writer_add(bw, kv.first.c_str(), kv.second->dtype, static_cast<int>(kv.second->dims.size()),
                                   kv.second->dims.data(), copy.data(), copy.size()) != 0) {
          dtfn_bundle_writer_abandon(bw);
          return -21;
        }
      }
      l.unlock();
      return dtfn_bundle_writer_finish(bw) == 0 ? 0 : -22;
    }
    case OP_SHUTDOWN:
      request_stop();
      return 0;
    default:
      return -99;
  }
}

void dtfn_server::reap_finished() {
  std::vector<std::thread> done;
  {
    std::lock_guard<std::mutex> l(conn_mu);
    for (long long id : finished) {
      auto it = conns.find(id);
      if (it == conns.end()) continue;
      done.push_back(std::move(it->second.th));
      conns.erase(it);
    }
    finished.clear();
  }
  for (auto& t : done)
    if (t.joinable()) t.join();
}

void dtfn_server::handle(long long id, int fd) {
  tune_socket(fd);
  std::string buf;   // only ever grows: resize() would zero-fill megabytes per request otherwise
  for (;;) {
    uint32_t len;
    if (!recv_all(fd, &len, 4)) break;
    if (len < 1 || len > (1u << 30)) break;
    // the op byte lands at offset 3 so that the body behind it starts 4-byte aligned
    if (buf.size() < static_cast<size_t>(len) + 3) buf.resize(static_cast<size_t>(len) + 3);
    if (!recv_all(fd, &buf[3], len)) break;
    if (drop.load() > 0 && drop.fetch_sub(1) > 0) break;  // fault injection: vanish mid-request
    Reader r(buf.data() + 4, len - 1);
    Writer w;
    const uint8_t op = static_cast<uint8_t>(buf[3]);
    int status = dispatch(op, r, w);
    if (!r.ok) status = -1;
    char hdr[8];
    const uint32_t olen = static_cast<uint32_t>(4 + w.s.size());
    const int32_t st = status;
    memcpy(hdr, &olen, 4);
    memcpy(hdr + 4, &st, 4);
    if (!send_frame(fd, hdr, 8, w.s)) break;
    if (op == OP_SHUTDOWN) break;
  }
  {
    // forget the descriptor before closing it: stop() must never shutdown() a number that the process
    // has meanwhile handed to another socket
    std::lock_guard<std::mutex> l(conn_mu);
    auto it = conns.find(id);
    if (it != conns.end()) it->second.fd = -1;
    finished.push_back(id);
  }
  ::shutdown(fd, SHUT_RDWR);
  ::close(fd);
}

extern "C" {

dtfn_server* dtfn_server_start(const char* host, int port) {
  const int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) {
    set_error("server: socket() failed: %s", strerror(errno));
    return nullptr;
  }
  int one = 1;
  setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_port = htons(static_cast<uint16_t>(port));
  if (host == nullptr || host[0] == '\0' || strcmp(host, "localhost") == 0) host = "127.0.0.1";
  if (strcmp(host, "0.0.0.0") == 0) {
    addr.sin_addr.s_addr = INADDR_ANY;
  } else if (inet_pton(AF_INET, host, &addr.sin_addr) != 1) {
    addr.sin_addr.s_addr = INADDR_ANY;  // named host: listen on every interface
  }
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
    set_error("server: bind to %s:%d failed: %s", host, port, strerror(errno));
    ::close(fd);
    return nullptr;
  }
  if (::listen(fd, 128) != 0) {
    set_error("server: listen failed: %s", strerror(errno));
    ::close(fd);
    return nullptr;
  }
  socklen_t alen = sizeof(addr);
  getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &alen);
  auto* s = new dtfn_server();
  s->listen_fd = fd;
  s->port = ntohs(addr.sin_port);
  s->accept_thread = std::thread([s] {
    while (!s->stopping) {
      pollfd p{s->listen_fd, POLLIN, 0};
      const int pr = ::poll(&p, 1, 100);
      if (pr <= 0) continue;
      const int c = ::accept(s->listen_fd, nullptr, nullptr);
      if (c < 0) continue;
      s->reap_finished();
      std::lock_guard<std::mutex> l(s->conn_mu);
      const long long id = s->next_conn_id++;
      auto& conn = s->conns[id];
      conn.fd = c;
      conn.th = std::thread([s, id, c] { s->handle(id, c); });
    }
  });
  return s;
}

int dtfn_server_port(const dtfn_server* s) { return s->port; }

void dtfn_server_join(dtfn_server* s) {
  std::unique_lock<std::mutex> l(s->join_mu);
  s->join_cv.wait(l, [s] { return s->stopped; });
}

void dtfn_server_stop(dtfn_server* s) {
  s->request_stop();
  if (s->accept_thread.joinable()) s->accept_thread.join();
  ::close(s->listen_fd);
  {
    std::lock_guard<std::mutex> l(s->conn_mu);
    for (auto& kv : s->conns)
      if (kv.second.fd >= 0) ::shutdown(kv.second.fd, SHUT_RDWR);
  }
  // wake any handler blocked in a queue wait
  {
    std::lock_guard<std::mutex> l(s->store.qmu);
    for (auto& kv : s->store.queues) kv.second->cv.notify_all();
  }
  std::vector<std::thread> all;
  {
    std::lock_guard<std::mutex> l(s->conn_mu);
    for (auto& kv : s->conns) all.push_back(std::move(kv.second.th));
  }
  for (auto& t : all)
    if (t.joinable()) t.join();
  delete s;
}

void dtfn_server_inject_drop(dtfn_server* s, int n) { s->drop = n; }

long long dtfn_server_num_variables(const dtfn_server* s) {
  auto* ms = const_cast<dtfn_server*>(s);
  std::shared_lock<std::shared_mutex> l(ms->store.mu);
  return static_cast<long long>(ms->store.vars.size());
}

}  // extern "C"

// ------------------------------------------------------------------------------- client
static int open_client_socket(const char* host, int port, int timeout_ms);

struct dtfn_client {
  int fd = -1;
  std::string host;
  int port = 0;
  int timeout_ms = 0;
  std::mutex mu;
  std::string resp;

  // A request that failed in transit may still be answered later (a slow SAVE, an ENQUEUE held on a
  // full queue): that late reply would be read as the answer to the next request.  So the socket of a
  // failed exchange is never reused -- it is closed, and the next call starts on a fresh connection.
  void poison() {
    if (fd >= 0) {
      ::shutdown(fd, SHUT_RDWR);
      ::close(fd);
    }
    fd = -1;
  }

  // sends one request and receives the response payload into `resp`; returns status or <= -100 on
  // transport failure
  int call(uint8_t op, const Writer& w) {
    char hdr[5];
    const uint32_t len = static_cast<uint32_t>(1 + w.s.size());
    memcpy(hdr, &len, 4);
    hdr[4] = static_cast<char>(op);
    if (fd < 0) {
      fd = open_client_socket(host.c_str(), port, timeout_ms);
      if (fd < 0) return -100;   // error text set by open_client_socket
    }
    if (!send_frame(fd, hdr, 5, w.s)) {
      set_error("client: send to %s:%d failed: %s", host.c_str(), port, strerror(errno));
      poison();
      return -100;
    }
    uint32_t rlen;
    if (!recv_all(fd, &rlen, 4) || rlen < 4 || rlen > (1u << 30)) {
      set_error("client: no response from %s:%d (peer closed or timed out)", host.c_str(), port);
      poison();
      return -101;
    }
    int32_t status;
    if (!recv_all(fd, &status, 4)) {
      set_error("client: truncated response from %s:%d", host.c_str(), port);
      poison();
      return -101;
    }
    resp.resize(rlen - 4);
    if (rlen > 4 && !recv_all(fd, &resp[0], rlen - 4)) {
      set_error("client: truncated response from %s:%d", host.c_str(), port);
      poison();
      return -101;
    }
    if (status < 0 && status > -100) set_error("client: server returned status %d for op %d", status, (int)op);
    return status;
  }
};

static void set_sock_timeout(int fd, int timeout_ms);
static int open_client_socket(const char* host, int port, int timeout_ms) {
  addrinfo hints{}, *res = nullptr;
  hints.ai_family = AF_INET;
  hints.ai_socktype = SOCK_STREAM;
  char portstr[16];
  snprintf(portstr, sizeof(portstr), "%d", port);
  if (getaddrinfo(host, portstr, &hints, &res) != 0 || !res) {
    set_error("client: cannot resolve %s", host);
    return -1;
  }
  const int fd = ::socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) {
    freeaddrinfo(res);
    set_error("client: socket() failed");
    return -1;
  }
  if (timeout_ms > 0) set_sock_timeout(fd, timeout_ms);
  const int rc = ::connect(fd, res->ai_addr, res->ai_addrlen);
  freeaddrinfo(res);
  if (rc != 0) {
    set_error("client: connect to %s:%d failed: %s", host, port, strerror(errno));
    ::close(fd);
    return -1;
  }
  tune_socket(fd);
  return fd;
}

static void set_sock_timeout(int fd, int timeout_ms) {
  timeval tv{};
  tv.tv_sec = timeout_ms / 1000;
  tv.tv_usec = (timeout_ms % 1000) * 1000;
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
}

extern "C" {

dtfn_client* dtfn_client_connect(const char* host, int port, int timeout_ms) {
  if (host == nullptr || host[0] == '\0' || strcmp(host, "localhost") == 0) host = "127.0.0.1";
  const int fd = open_client_socket(host, port, timeout_ms);
  if (fd < 0) return nullptr;
  auto* c = new dtfn_client();
  c->fd = fd;
  c->host = host;
  c->port = port;
  c->timeout_ms = timeout_ms;
  return c;
}

void dtfn_client_close(dtfn_client* c) {
  if (!c) return;
  c->poison();
  delete c;
}

void dtfn_client_set_timeout(dtfn_client* c, int timeout_ms) {
  std::lock_guard<std::mutex> l(c->mu);
  c->timeout_ms = timeout_ms;
  if (c->fd >= 0) set_sock_timeout(c->fd, timeout_ms);
}

int dtfn_client_ping(dtfn_client* c) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  return c->call(OP_PING, w);
}

int dtfn_client_init_var(dtfn_client* c, const char* name, int dtype, int ndims, const long long* dims,
                         const void* data, size_t nbytes, int overwrite) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  w.i32(dtype);
  w.i32(ndims);
  for (int i = 0; i < ndims; ++i) w.i64(dims[i]);
  w.u8(static_cast<uint8_t>(overwrite));
  w.bytes(data, nbytes);
  return c->call(OP_INIT_VAR, w);
}

long long dtfn_client_pull(dtfn_client* c, const char* name, void* dst, size_t cap) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  const int st = c->call(OP_PULL, w);
  if (st != 0) return st;
  Reader r(c->resp.data(), c->resp.size());
  size_t n;
  const char* p = r.bytes(&n);
  if (!r.ok) return -102;
  if (n > cap) {
    set_error("client: variable '%s' needs %zu bytes, buffer has %zu", name, n, cap);
    return -103;
  }
  memcpy(dst, p, n);
  return static_cast<long long>(n);
}

long long dtfn_client_list(dtfn_client* c, char* buf, size_t cap) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  const int st = c->call(OP_LIST, w);
  if (st != 0) return st;
  Reader r(c->resp.data(), c->resp.size());
  const std::string s = r.str();
  if (!r.ok) return -102;
  if (buf && cap > 0) {
    const size_t n = std::min(cap - 1, s.size());
    memcpy(buf, s.data(), n);
    buf[n] = '\0';
  }
  return static_cast<long long>(s.size());
}

int dtfn_client_var_info(dtfn_client* c, const char* name, int* dtype, long long* dims, long long* nbytes) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  const int st = c->call(OP_VAR_INFO, w);
  if (st != 0) return st < 0 ? st : -st;
  Reader r(c->resp.data(), c->resp.size());
  *dtype = r.i32();
  const int nd = r.i32();
  for (int i = 0; i < nd && i < 8; ++i) dims[i] = r.i64();
  *nbytes = r.i64();
  return r.ok ? nd : -102;
}

int dtfn_client_push_apply(dtfn_client* c, const char* name, int optimizer, float lr, const float* grad, size_t n) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  w.i32(optimizer);
  w.f32(lr);
  w.bytes(grad, n * 4);
  return c->call(OP_PUSH_APPLY, w);
}

int dtfn_client_push_pull(dtfn_client* c, int k, const char* const* names, int optimizer, float lr,
                          const float* const* grads, const size_t* lens, const char* step_name,
                          long long step_inc, float* const* dst, long long* step_out) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  size_t total = 64;
  for (int i = 0; i < k; ++i) total += lens[i] * 4 + strlen(names[i]) + 16;
  w.s.reserve(total);
  w.i32(k);
  w.i32(optimizer);
  w.f32(lr);
  for (int i = 0; i < k; ++i) {
    w.str(names[i]);
    w.bytes(grads[i], lens[i] * 4);
  }
  w.str(step_name ? step_name : "");
  w.i64(step_inc);
  w.u8(dst != nullptr ? 1 : 0);
  w.i32(0);
  const int st = c->call(OP_PUSH_PULL, w);
  if (st != 0) return st;
  Reader r(c->resp.data(), c->resp.size());
  const int64_t step = r.i64();
  if (step_out) *step_out = step;
  if (dst) {
    for (int i = 0; i < k; ++i) {
      size_t n;
      const char* p = r.bytes(&n);
      if (!r.ok) return -102;
      if (n != lens[i] * 4) {
        set_error("client: pulled '%s' has %zu bytes, expected %zu", names[i], n, lens[i] * 4);
        return -103;
      }
      memcpy(dst[i], p, n);
    }
  }
  return r.ok ? 0 : -102;
}

int dtfn_client_assign_add_i64(dtfn_client* c, const char* name, long long delta, long long* out) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  w.i64(delta);
  const int st = c->call(OP_ASSIGN_ADD, w);
  if (st != 0) return st;
  Reader r(c->resp.data(), c->resp.size());
  const int64_t v = r.i64();
  if (out) *out = v;
  return r.ok ? 0 : -102;
}

int dtfn_client_queue_create(dtfn_client* c, const char* name, int capacity) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(name);
  w.i32(capacity);
  return c->call(OP_QUEUE_CREATE, w);
}

int dtfn_client_enqueue(dtfn_client* c, const char* name, int value) {
  // a full bounded queue holds the request for at most `slice` ms per attempt (below the socket
  // timeout); -2 = still full, ask again, up to the connection's timeout over all attempts
  const int budget_ms = c->timeout_ms > 0 ? c->timeout_ms : 30000;
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(budget_ms);
  for (;;) {
    const auto left =
        std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
    if (left <= 0) {
      set_error("client: queue '%s' on %s:%d stayed full for %d ms", name, c->host.c_str(), c->port, budget_ms);
      return -2;
    }
    std::lock_guard<std::mutex> l(c->mu);
    Writer w;
    w.str(name);
    w.i32(value);
    w.i32(static_cast<int>(std::min<long long>(500, left)));
    const int st = c->call(OP_ENQUEUE, w);
    if (st != -2) return st;
  }
}

int dtfn_client_dequeue(dtfn_client* c, const char* name, int* value, int timeout_ms) {
  // the server waits in slices of <= 500 ms so that socket timeouts and shutdown stay responsive
  const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
  for (;;) {
    int slice = 500;
    if (timeout_ms > 0) {
      const auto left =
          std::chrono::duration_cast<std::chrono::milliseconds>(deadline - std::chrono::steady_clock::now()).count();
      if (left <= 0) return -2;
      slice = static_cast<int>(std::min<long long>(500, left));
    }
    std::lock_guard<std::mutex> l(c->mu);
    Writer w;
    w.str(name);
    w.i32(slice);
    const int st = c->call(OP_DEQUEUE, w);
    if (st == 0) {
      Reader r(c->resp.data(), c->resp.size());
      *value = r.i32();
      return r.ok ? 0 : -102;
    }
    if (st != -2) return st;
  }
}

int dtfn_client_save(dtfn_client* c, const char* prefix, const char* exclude_newline_joined) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  w.str(prefix);
  std::vector<std::string> excl;
  if (exclude_newline_joined && exclude_newline_joined[0]) {
    std::string all(exclude_newline_joined), cur;
    for (char ch : all) {
      if (ch == '\n') { if (!cur.empty()) excl.push_back(cur); cur.clear(); }
      else cur.push_back(ch);
    }
    if (!cur.empty()) excl.push_back(cur);
  }
  w.i32(static_cast<int32_t>(excl.size()));
  for (auto& e : excl) w.str(e);
  return c->call(OP_SAVE, w);
}

int dtfn_client_shutdown_server(dtfn_client* c) {
  std::lock_guard<std::mutex> l(c->mu);
  Writer w;
  return c->call(OP_SHUTDOWN, w);
}

}  // extern "C"
