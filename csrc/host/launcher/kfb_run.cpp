// kfb-run: start N ranks of one command on this machine, one per GPU.
//
//   kfb-run -np N [-port P] [-timeout SECONDS] [-logdir DIR] [-grace SECONDS] [--] cmd [args...]
//
// The role `kungfu-run -np N` (a Go binary of the external KungFu runtime) and `mpirun -np N`
// play in the reference's launch scripts (T/README.md:97-114, T/run_kf.sh-bak:18-21). Every rank
// gets RANK / LOCAL_RANK / WORLD_SIZE / LOCAL_WORLD_SIZE / MASTER_ADDR=127.0.0.1 / MASTER_PORT and
// its own process group. Rank 0's output passes through; other ranks' lines are prefixed with
// "[rank r] ". With -logdir every rank's stdout / stderr is also written to
// DIR/127.0.0.1.<10000+r>.{stdout,stderr}.log, the naming kungfu-run left in the reference tree
// (T/127.0.0.1.1000*.std*.log). The first rank that fails ends the job: the launcher reports
// it, sends SIGTERM to every other rank's process group and SIGKILL after the grace period -
// the failure mode recorded in the reference's logs ("exit on error: N tasks failed", the rest
// "signal: killed"; T/slurm-3032311.out:168-171). SIGINT / SIGTERM sent to the
// launcher are forwarded. Exit status: 0, the first failing rank's status (128 + signal when it
// was killed), or 124 on timeout.
#include <arpa/inet.h>
#include <errno.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <poll.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

volatile sig_atomic_t g_signal = 0;
void on_signal(int sig) { g_signal = sig; }

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

int free_port() {
  int fd = socket(AF_INET, SOCK_STREAM, 0);
  if (fd < 0) return 29500;
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  addr.sin_port = 0;
  int port = 29500;
  socklen_t len = sizeof(addr);
  if (bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) == 0 &&
      getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len) == 0)
    port = ntohs(addr.sin_port);
  close(fd);
  return port;
}

struct Stream {   // one pipe of one rank
  int fd = -1;
  int rank = 0;
  bool is_err = false;
  FILE* log = nullptr;
  std::string partial;
};

struct Rank {
  pid_t pid = -1;
  bool done = false;
  int status = 0;
};

void emit(const Stream& s, const char* data, size_t n) {
  FILE* out = s.is_err ? stderr : stdout;
  if (s.rank != 0) std::fprintf(out, "[rank %d] ", s.rank);
  std::fwrite(data, 1, n, out);
  std::fputc('\n', out);
  std::fflush(out);
  if (s.log != nullptr) {
    std::fwrite(data, 1, n, s.log);
    std::fputc('\n', s.log);
    std::fflush(s.log);
  }
}

// forwards whatever is readable; returns false at end of file
bool pump(Stream& s) {
  char buf[4096];
  const ssize_t got = read(s.fd, buf, sizeof(buf));
  if (got < 0) return errno == EAGAIN || errno == EINTR;
  if (got == 0) {
    if (!s.partial.empty()) emit(s, s.partial.data(), s.partial.size());
    s.partial.clear();
    return false;
  }
  s.partial.append(buf, static_cast<size_t>(got));
  size_t start = 0;
  for (;;) {
    const size_t nl = s.partial.find('\n', start);
    if (nl == std::string::npos) break;
    emit(s, s.partial.data() + start, nl - start);
    start = nl + 1;
  }
  s.partial.erase(0, start);
  return true;
}

void signal_all(const std::vector<Rank>& ranks, int sig) {
  for (const Rank& r : ranks)
    if (!r.done && r.pid > 0) kill(-r.pid, sig);  // the rank's whole process group
}

int exit_code_of(int status) {
  if (WIFEXITED(status)) return WEXITSTATUS(status);
  if (WIFSIGNALED(status)) return 128 + WTERMSIG(status);
  return 1;
}

int usage() {
  std::fprintf(stderr,
               "usage: kfb-run -np N [-port P] [-timeout SECONDS] [-logdir DIR] [-grace SECONDS] "
               "[--] command [args...]\n");
  return 2;
}

}  // namespace

int main(int argc, char** argv) {
  int np = 0, port = 0, i = 1;
  double timeout = 0.0, grace = 5.0;
  std::string logdir;
  for (; i < argc; ++i) {
    const std::string a = argv[i];
    auto need = [&](const char* what) -> const char* {
      if (i + 1 >= argc) {
        std::fprintf(stderr, "kfb-run: %s needs a value\n", what);
        std::exit(2);
      }
      return argv[++i];
    };
    if (a == "-np") {
      np = std::atoi(need("-np"));
    } else if (a == "-port") {
      port = std::atoi(need("-port"));
    } else if (a == "-timeout") {
      timeout = std::atof(need("-timeout"));
    } else if (a == "-grace") {
      grace = std::atof(need("-grace"));
    } else if (a == "-logdir") {
      logdir = need("-logdir");
    } else if (a == "--") {
      ++i;
      break;
    } else {
      break;  // first word of the command
    }
  }
  if (np < 1 || np > 1024 || i >= argc) return usage();
  if (port <= 0) port = free_port();
  if (!logdir.empty()) mkdir(logdir.c_str(), 0777);

  struct sigaction sa {};
  sa.sa_handler = on_signal;
  sigaction(SIGINT, &sa, nullptr);
  sigaction(SIGTERM, &sa, nullptr);
  signal(SIGPIPE, SIG_IGN);

  std::vector<Rank> ranks(np);
  std::vector<Stream> streams;
  for (int r = 0; r < np; ++r) {
    int out_pipe[2], err_pipe[2];
    if (pipe(out_pipe) != 0 || pipe(err_pipe) != 0) {
      std::perror("kfb-run: pipe");
      signal_all(ranks, SIGKILL);
      return 1;
    }
    const pid_t pid = fork();
    if (pid < 0) {
      std::perror("kfb-run: fork");
      signal_all(ranks, SIGKILL);
      return 1;
    }
    if (pid == 0) {
      setsid();  // own session / process group: the launcher can stop everything the rank started
      dup2(out_pipe[1], STDOUT_FILENO);
      dup2(err_pipe[1], STDERR_FILENO);
      close(out_pipe[0]);
      close(out_pipe[1]);
      close(err_pipe[0]);
      close(err_pipe[1]);
      for (const Stream& s : streams) close(s.fd);
      const std::string rs = std::to_string(r), ws = std::to_string(np), ps = std::to_string(port);
      setenv("RANK", rs.c_str(), 1);
      setenv("LOCAL_RANK", rs.c_str(), 1);
      setenv("WORLD_SIZE", ws.c_str(), 1);
      setenv("LOCAL_WORLD_SIZE", ws.c_str(), 1);
      setenv("MASTER_ADDR", "127.0.0.1", 1);
      setenv("MASTER_PORT", ps.c_str(), 1);
      execvp(argv[i], argv + i);
      std::fprintf(stderr, "kfb-run: cannot execute %s: %s\n", argv[i], std::strerror(errno));
      _exit(127);
    }
    close(out_pipe[1]);
    close(err_pipe[1]);
    ranks[r].pid = pid;
    for (int e = 0; e < 2; ++e) {
      Stream s;
      s.fd = e ? err_pipe[0] : out_pipe[0];
      s.rank = r;
      s.is_err = e != 0;
      fcntl(s.fd, F_SETFL, fcntl(s.fd, F_GETFL) | O_NONBLOCK);
      fcntl(s.fd, F_SETFD, FD_CLOEXEC);
      if (!logdir.empty()) {
        const std::string path = logdir + "/127.0.0.1." + std::to_string(10000 + r) +
                                 (e ? ".stderr.log" : ".stdout.log");
        s.log = std::fopen(path.c_str(), "w");
      }
      streams.push_back(s);
    }
  }

  const double t0 = now_s();
  int rc = 0, live = np;
  bool stopping = false, timed_out = false;
  double kill_at = 0.0;
  auto stop_all = [&](const char* why) {
    if (stopping) return;
    stopping = true;
    kill_at = now_s() + grace;
    std::fprintf(stderr, "[kfb-run] %s; stopping the other ranks\n", why);
    signal_all(ranks, SIGTERM);
  };
  while (live > 0 || !streams.empty()) {
    std::vector<pollfd> fds;
    for (const Stream& s : streams) fds.push_back({s.fd, POLLIN, 0});
    poll(fds.data(), fds.size(), 50);
    for (size_t k = 0; k < streams.size();) {
      if ((fds[k].revents & (POLLIN | POLLHUP | POLLERR)) && !pump(streams[k])) {
        close(streams[k].fd);
        if (streams[k].log != nullptr) std::fclose(streams[k].log);
        streams.erase(streams.begin() + k);
        fds.erase(fds.begin() + k);
      } else {
        ++k;
      }
    }
    for (int r = 0; r < np; ++r) {
      if (ranks[r].done) continue;
      int status = 0;
      const pid_t w = waitpid(ranks[r].pid, &status, WNOHANG);
      if (w != ranks[r].pid) continue;
      ranks[r].done = true;
      ranks[r].status = status;
      --live;
      const int code = exit_code_of(status);
      if (code != 0 && rc == 0 && !stopping) {
        rc = code;
        char why[128];
        if (WIFSIGNALED(status))
          std::snprintf(why, sizeof(why), "rank %d was killed by signal %d", r, WTERMSIG(status));
        else
          std::snprintf(why, sizeof(why), "rank %d exited with %d", r, code);
        stop_all(why);
      }
    }
    if (g_signal != 0 && !stopping) {
      rc = 128 + g_signal;
      stop_all("interrupted");
    }
    if (timeout > 0 && !stopping && now_s() - t0 > timeout) {
      timed_out = true;
      char why[64];
      std::snprintf(why, sizeof(why), "timeout after %.0f s", timeout);
      stop_all(why);
    }
    if (stopping && kill_at > 0 && now_s() > kill_at) {
      signal_all(ranks, SIGKILL);
      kill_at = 0.0;
    }
    if (live == 0 && stopping) {
      // a killed rank's grandchildren may still hold the pipes open: do not wait for them
      for (Stream& s : streams) {
        pump(s);
        close(s.fd);
        if (s.log != nullptr) std::fclose(s.log);
      }
      streams.clear();
    }
  }
  if (timed_out) return 124;
  return rc;
}
