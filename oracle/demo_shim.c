/* TEST INFRASTRUCTURE. Linked into oracle/_ref/field_d_planner_1_no_heur (the reference's demo process, compiled from
 * /root/reference/Tests/Planners/FDSTAR/main.cpp as it lies). main.cpp:33-40 pins the process to CPU 0 and abort()s if the
 * call fails; whether CPU 0 is in the container's cpuset has nothing to do with the planner, so the call is answered here
 * (SURVEY.md App. B-10). Nothing else of the process is touched. */
#define _GNU_SOURCE
#include <pthread.h>
#include <sched.h>
#include <stddef.h>

int pthread_setaffinity_np(pthread_t th, size_t cpusetsize, const cpu_set_t *cpuset) {
    (void) th; (void) cpusetsize; (void) cpuset;
    return 0;
}
