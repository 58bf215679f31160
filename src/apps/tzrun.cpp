// tzrun -n N [--port P] [--grace S] prog [args...]: start N ranks of a native driver on this node
// (one per GPU) with RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT set, the
// way mpirun/srun start the reference's examples (docs/perlmutter.md:23-24).
//
//  * SIGINT / SIGTERM / SIGABRT received by tzrun are forwarded to every rank, so a batch
//    system's "--signal=SIGABRT@10" still makes rank 0 dump its partial results
//    (tenzing::register_handler; reference: scripts/perlmutter/spmv.sh:12).
//  * If a rank dies (non-zero exit or a signal) while others are still running, the survivors
//    would wait for it forever in their next collective (in the reference "a rank failure hangs
//    the job").  tzrun gives them S seconds (default 10) to finish on their own, then sends
//    SIGTERM and, five seconds later, SIGKILL.
// Exit code: the first non-zero rank exit code (128 + signal number for a signalled rank).
#include <cerrno>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include <sys/wait.h>
#include <unistd.h>

static volatile sig_atomic_t g_forward = 0; // last signal to pass on to the ranks

static void on_signal(int sig) { g_forward = sig; }

static double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec);
}

int main(int argc, char **argv) {
  int n = 1, port = 29400, i = 1;
  double grace = 10.0;
  for (; i < argc; ++i) {
    if (!std::strcmp(argv[i], "-n") && i + 1 < argc) n = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--port") && i + 1 < argc) port = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--grace") && i + 1 < argc) grace = std::atof(argv[++i]);
    else break;
  }
  if (i >= argc || n < 1) {
    std::fprintf(stderr, "usage: tzrun -n N [--port P] [--grace S] prog [args...]\n");
    return 64;
  }
  std::vector<pid_t> pids;
  for (int r = 0; r < n; ++r) {
    pid_t pid = fork();
    if (pid < 0) {
      std::perror("fork");
      for (pid_t p : pids) kill(p, SIGTERM);
      return 71;
    }
    if (pid == 0) {
      setenv("RANK", std::to_string(r).c_str(), 1);
      setenv("LOCAL_RANK", std::to_string(r).c_str(), 1);
      setenv("WORLD_SIZE", std::to_string(n).c_str(), 1);
      setenv("MASTER_ADDR", "127.0.0.1", 0);
      setenv("MASTER_PORT", std::to_string(port).c_str(), 1);
      execvp(argv[i], argv + i);
      std::perror("execvp");
      _exit(127);
    }
    pids.push_back(pid);
  }

  struct sigaction sa;
  std::memset(&sa, 0, sizeof(sa));
  sa.sa_handler = on_signal;
  sigemptyset(&sa.sa_mask);
  for (int sig : {SIGINT, SIGTERM, SIGABRT}) sigaction(sig, &sa, nullptr);

  std::vector<bool> alive(pids.size(), true);
  size_t nAlive = pids.size();
  int rc = 0;
  double failedAt = -1; // when the first rank died abnormally
  bool termSent = false, killSent = false;
  auto signal_all = [&](int sig) {
    for (size_t r = 0; r < pids.size(); ++r)
      if (alive[r]) kill(pids[r], sig);
  };
  while (nAlive) {
    if (g_forward) {
      const int sig = g_forward;
      g_forward = 0;
      signal_all(sig);
    }
    bool reaped = false;
    for (size_t r = 0; r < pids.size(); ++r) {
      if (!alive[r]) continue;
      int st = 0;
      const pid_t got = waitpid(pids[r], &st, WNOHANG);
      if (got == 0 || (got < 0 && errno == EINTR)) continue;
      alive[r] = false;
      --nAlive;
      reaped = true;
      const int code = got < 0 ? 1 : (WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st));
      if (code) {
        if (!rc) rc = code;
        if (failedAt < 0 && nAlive) {
          failedAt = now_s();
          std::fprintf(stderr, "tzrun: rank %zu ended with code %d; the other ranks get %.0f s\n", r,
                       code, grace);
        }
      }
    }
    if (failedAt >= 0 && nAlive) {
      const double waited = now_s() - failedAt;
      if (!termSent && waited >= grace) {
        std::fprintf(stderr, "tzrun: terminating %zu remaining rank(s)\n", nAlive);
        signal_all(SIGTERM);
        termSent = true;
      } else if (termSent && !killSent && waited >= grace + 5.0) {
        signal_all(SIGKILL);
        killSent = true;
      }
    }
    if (!reaped && nAlive) usleep(20000);
  }
  return rc;
}
