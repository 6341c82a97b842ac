// ThreadSanitizer stress of the C++ parameter-server transport.
//
//   g++ -std=c++17 -O1 -g -fsanitize=thread -pthread -I include csrc/native/*.cc -o /tmp/dtf_tsan_stress
//   TSAN_OPTIONS="suppressions=csrc/native/tsan.supp halt_on_error=1" /tmp/dtf_tsan_stress
//
// One in-process server, several client threads: variable creation races, fused push_pull steps,
// a shared-name queue, listing, a ps-side checkpoint save while training runs, fault-injected dropped
// requests and shutdown.  The optimizer applies and the copy-out of variable values run without locks
// on purpose (Hogwild, TF's use_locking=False); those two code paths -- and only those -- are listed
// in tsan.supp.  Anything else ThreadSanitizer reports is a bug.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "dtf_native.h"

#define CHECK(cond)                                                               \
  do {                                                                            \
    if (!(cond)) {                                                                \
      fprintf(stderr, "CHECK failed %s:%d: %s (%s)\n", __FILE__, __LINE__, #cond, dtfn_last_error()); \
      exit(2);                                                                    \
    }                                                                             \
  } while (0)

int main(int argc, char** argv) {
  const int kThreads = 6, kSteps = argc > 1 ? atoi(argv[1]) : 300;
  const char* tmp = argc > 2 ? argv[2] : "/tmp";
  dtfn_server* srv = dtfn_server_start("127.0.0.1", 0);
  CHECK(srv != nullptr);
  const int port = dtfn_server_port(srv);
  std::atomic<long long> steps_done{0};
  std::vector<std::thread> ts;
  for (int t = 0; t < kThreads; ++t) {
    ts.emplace_back([&, t] {
      dtfn_client* c = dtfn_client_connect("127.0.0.1", port, 5000);
      CHECK(c != nullptr);
      CHECK(dtfn_client_ping(c) == 0);
      // every thread tries to create the same variables: exactly one wins, nobody crashes
      const long long dims[1] = {64};
      std::vector<float> init(64, 1.0f);
      const long long zero = 0, sdims[1] = {1};
      CHECK(dtfn_client_init_var(c, "w", 1, 1, dims, init.data(), init.size() * 4, 0) >= 0);
      CHECK(dtfn_client_init_var(c, "b", 1, 1, dims, init.data(), init.size() * 4, 0) >= 0);
      CHECK(dtfn_client_init_var(c, "global_step", 9, 0, sdims, &zero, 8, 0) >= 0);
      CHECK(dtfn_client_queue_create(c, "done", 64) == 0);
      std::vector<float> g(64, 0.01f), w(64), b(64);
      const char* names[2] = {"w", "b"};
      const float* grads[2] = {g.data(), g.data()};
      const size_t lens[2] = {64, 64};
      float* dst[2] = {w.data(), b.data()};
      for (int s = 0; s < kSteps; ++s) {
        long long step = 0;
        const int opt = (t % 3 == 0) ? 0 : ((t % 3 == 1) ? 2 : 3);   // sgd / adagrad / adam side by side
        CHECK(dtfn_client_push_pull(c, 2, names, opt, 0.01f, grads, lens, "global_step", 1, dst, &step) == 0);
        CHECK(step >= 1);
        if (s % 50 == 0) {
          char buf[4096];
          CHECK(dtfn_client_list(c, buf, sizeof(buf)) > 0);
          CHECK(dtfn_client_pull(c, "w", w.data(), 256) == 256);
        }
        if (t == 0 && s % 100 == 50) {
          const std::string prefix = std::string(tmp) + "/tsan_ckpt_" + std::to_string(s);
          CHECK(dtfn_client_save(c, prefix.c_str(), nullptr) == 0);    // snapshot while others train
        }
        steps_done.fetch_add(1);
      }
      CHECK(dtfn_client_enqueue(c, "done", t) == 0);
      dtfn_client_close(c);
    });
  }
  // the "ps wait loop": one more client drains the queue
  {
    dtfn_client* c = dtfn_client_connect("127.0.0.1", port, 5000);
    CHECK(c != nullptr);
    CHECK(dtfn_client_queue_create(c, "done", 64) == 0);
    for (int i = 0; i < kThreads; ++i) {
      int v = -1, rc;
      do {
        rc = dtfn_client_dequeue(c, "done", &v, 200);
      } while (rc == -2);
      CHECK(rc == 0 && v >= 0 && v < kThreads);
    }
    long long step = 0;
    CHECK(dtfn_client_assign_add_i64(c, "global_step", 0, &step) == 0);
    CHECK(step == static_cast<long long>(kThreads) * kSteps);
    // fault injection: the next request vanishes, the client sees an error instead of hanging
    dtfn_client_set_timeout(c, 300);
    dtfn_server_inject_drop(srv, 1);
    CHECK(dtfn_client_ping(c) != 0);
    dtfn_client_close(c);
  }
  for (auto& t : ts) t.join();
  CHECK(dtfn_server_num_variables(srv) >= 3);
  dtfn_server_stop(srv);
  printf("tsan_stress ok: %lld steps from %d threads\n", steps_done.load(), kThreads);
  return 0;
}
