// tzrun -n N [--port P] prog [args...]: start N ranks of a native driver on this node
// (one per GPU) with RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT set, the
// way mpirun/srun start the reference's examples (docs/perlmutter.md:23-24).  Exit code is
// the first non-zero rank exit code.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <sys/wait.h>
#include <unistd.h>

int main(int argc, char **argv) {
  int n = 1, port = 29400, i = 1;
  for (; i < argc; ++i) {
    if (!std::strcmp(argv[i], "-n") && i + 1 < argc) n = std::atoi(argv[++i]);
    else if (!std::strcmp(argv[i], "--port") && i + 1 < argc) port = std::atoi(argv[++i]);
    else break;
  }
  if (i >= argc || n < 1) {
    std::fprintf(stderr, "usage: tzrun -n N [--port P] prog [args...]\n");
    return 64;
  }
  std::vector<pid_t> pids;
  for (int r = 0; r < n; ++r) {
    pid_t pid = fork();
    if (pid < 0) {
      std::perror("fork");
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
  int rc = 0;
  for (pid_t pid : pids) {
    int st = 0;
    waitpid(pid, &st, 0);
    const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
    if (code && !rc) rc = code;
  }
  return rc;
}
