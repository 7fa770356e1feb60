// jb_mpirun -np P prog args... : fork/exec launcher of the in-tree MPI shim.
// Creates one AF_UNIX stream socketpair per unordered rank pair, forks P children, hands each its
// P-1 descriptors through JB_MPI_FDS, and returns the largest child exit code (like mpirun, the job
// fails if any rank fails; the reference returns 1 / 2 from every rank, main.cpp:82,89).
#include <errno.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

int main(int argc, char** argv) {
  int np = 1, i = 1;
  std::vector<std::pair<std::string, std::string>> rank_env;  // --rank-env NAME : NAME=<rank>
  while (i < argc && argv[i][0] == '-') {
    if ((!strcmp(argv[i], "-np") || !strcmp(argv[i], "-n")) && i + 1 < argc) {
      np = atoi(argv[i + 1]);
      i += 2;
    } else if (!strcmp(argv[i], "--rank-env") && i + 1 < argc) {
      rank_env.push_back({argv[i + 1], ""});
      i += 2;
    } else {
      break;
    }
  }
  if (np < 1 || i >= argc) {
    fprintf(stderr, "usage: %s -np P [--rank-env NAME] prog [args...]\n", argv[0]);
    return 64;
  }
  std::vector<std::vector<int>> fd(np, std::vector<int>(np, -1));
  const int bufsz = 4 << 20;
  for (int a = 0; a < np; ++a)
    for (int b = a + 1; b < np; ++b) {
      int sv[2];
      if (socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) {
        perror("socketpair");
        return 71;
      }
      for (int k = 0; k < 2; ++k) {
        setsockopt(sv[k], SOL_SOCKET, SO_SNDBUF, &bufsz, sizeof(bufsz));
        setsockopt(sv[k], SOL_SOCKET, SO_RCVBUF, &bufsz, sizeof(bufsz));
      }
      fd[a][b] = sv[0];
      fd[b][a] = sv[1];
    }
  std::vector<pid_t> pids(np);
  for (int r = 0; r < np; ++r) {
    pid_t pid = fork();
    if (pid < 0) {
      perror("fork");
      return 71;
    }
    if (pid == 0) {
      std::string fds;
      for (int q = 0; q < np; ++q) {
        if (q) fds += ",";
        fds += std::to_string(fd[r][q]);
      }
      for (int a = 0; a < np; ++a)
        for (int b = 0; b < np; ++b)
          if (a != r && fd[a][b] >= 0) close(fd[a][b]);
      setenv("JB_MPI_RANK", std::to_string(r).c_str(), 1);
      setenv("JB_MPI_SIZE", std::to_string(np).c_str(), 1);
      setenv("JB_MPI_FDS", fds.c_str(), 1);
      for (auto& e : rank_env) setenv(e.first.c_str(), std::to_string(r).c_str(), 1);
      execvp(argv[i], argv + i);
      fprintf(stderr, "jb_mpirun: cannot exec %s: %s\n", argv[i], strerror(errno));
      _exit(127);
    }
    pids[r] = pid;
  }
  for (int a = 0; a < np; ++a)
    for (int b = 0; b < np; ++b)
      if (fd[a][b] >= 0) close(fd[a][b]);
  int worst = 0;
  for (int r = 0; r < np; ++r) {
    int st = 0;
    if (waitpid(pids[r], &st, 0) < 0) continue;
    int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
    if (code > worst) worst = code;
    if (code >= 70) {  // a rank crashed: do not leave the others blocked in recv forever
      for (int q = 0; q < np; ++q)
        if (q != r) kill(pids[q], SIGTERM);
    }
  }
  return worst;
}
