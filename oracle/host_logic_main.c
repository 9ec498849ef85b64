/* TEST HARNESS only: runs the binding's host-logic scenario (ref_harness.c: ref_hook_host_logic) as a stand-alone program,
 * so that it can be built with -fsanitize=thread (make -C oracle tsan): the hook's once-per-picture / once-per-pair
 * protocols are exactly the kind of code a race detector should see. Usage: host_logic_tsan <path of the stand-in engine> */
#include <stdint.h>
#include <stdio.h>
int ref_hook_host_logic(const char *fake_library_path, uint64_t *report);
int main(int argc, char **argv) {
    uint64_t  r[9] = {0};
    const int rc = argc > 1 ? ref_hook_host_logic(argv[1], r) : -99;
    printf("ref_hook_host_logic rc %d; registers %llu releases %llu submits %llu waits %llu wait_pairs %llu\n", rc, (unsigned long long)r[0],
           (unsigned long long)r[1], (unsigned long long)r[2], (unsigned long long)r[3], (unsigned long long)r[4]);
    return rc ? 1 : 0;
}
