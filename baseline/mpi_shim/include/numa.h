/* Stand-in for libnuma's <numa.h>, which this image does not have.  The reference includes it
 * unconditionally (src/numa.cpp:8) but compiles every call out (its guard macro is never
 * defined), so declarations are all that is needed. */
#ifndef TZ_NUMA_STUB_H
#define TZ_NUMA_STUB_H
#ifdef __cplusplus
extern "C" {
#endif
struct bitmask {
  unsigned long size;
  unsigned long *maskp;
};
extern int numa_exit_on_warn, numa_exit_on_error;
extern struct bitmask *numa_all_nodes_ptr;
int numa_available(void);
void numa_set_strict(int);
void numa_set_bind_policy(int);
int numa_preferred(void);
int numa_max_node(void);
int numa_num_configured_nodes(void);
void numa_bind(struct bitmask *);
void numa_set_membind(struct bitmask *);
struct bitmask *numa_get_membind(void);
struct bitmask *numa_allocate_nodemask(void);
struct bitmask *numa_bitmask_setbit(struct bitmask *, unsigned int);
int numa_bitmask_isbitset(const struct bitmask *, unsigned int);
void numa_bitmask_free(struct bitmask *);
int numa_node_of_cpu(int);
#ifdef __cplusplus
}
#endif
#endif
