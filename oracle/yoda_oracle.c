/*
 * yoda_oracle.c -- CPU restatement of Yoda's Filter / CollectMaxValues / Score /
 * NormalizeScore arithmetic.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for the CUDA path.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call it.  The
 * product library (yoda_scheduler_b200/csrc) never links or calls anything in here.
 *
 * PARITY UNPINNED: the reference ships no tests, fixtures or golden vectors, and its Go
 * code cannot be compiled in this image (no Go toolchain, k8s/controller-runtime/SCV
 * modules not vendored).  The known-answer vectors in tests/golden/ were derived by hand
 * from the reference source semantics; a second, independent restatement
 * (oracle/pyref.py, arbitrary-precision ints) cross-checks this file.
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference).
 * Data layout is deliberately the reference's: array-of-structs Scv objects with string
 * Health fields, pod labels as a string map, Go `uint` = uint64 (GOARCH=amd64, Makefile:4).
 * Nothing here is copied; it is the same arithmetic re-expressed in C.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORA_OK 0
#define ORA_ERR_DIV0 (-2)   /* Go would panic: TotalMemorySum == 0 (algorithm.go:71,86) */
#define ORA_ERR_ARG (-1)

/* ------------------------------------------------------------------------------------------
 * [3P Go stdlib] strconv.Atoi, as used at filter.go:61,69,77.  Returns 1 on success.
 * Optional single leading '+'/'-', then one or more ASCII decimal digits, nothing else; the
 * value must fit int64, otherwise Go reports ErrRange and the callers below turn any error
 * into 0.
 * ---------------------------------------------------------------------------------------- */
static int go_atoi(const char *s, int64_t *out) {
    size_t n = strlen(s), i = 0;
    int neg = 0;
    if (n == 0) return 0;
    if (s[0] == '+' || s[0] == '-') {
        neg = (s[0] == '-');
        i = 1;
        if (n == 1) return 0;
    }
    /* magnitude limit: 2^63-1 for positive, 2^63 for negative */
    const uint64_t lim = neg ? ((uint64_t)1 << 63) : (((uint64_t)1 << 63) - 1);
    uint64_t acc = 0;
    for (; i < n; ++i) {
        unsigned d = (unsigned char)s[i] - (unsigned)'0';
        if (d > 9) return 0;
        if (acc > (lim - d) / 10) return 0; /* acc*10+d would exceed lim -> ErrRange */
        acc = acc * 10 + d;
    }
    *out = neg ? (int64_t)(0 - acc) : (int64_t)acc;
    return 1;
}

/* filter.go:60-66 strToUint, :68-74 StrToUint64 (identical on amd64), :76-82 StrToInt64 */
static uint64_t str_to_uint64(const char *s) {
    int64_t v;
    return go_atoi(s, &v) ? (uint64_t)v : 0;
}
static int64_t str_to_int64(const char *s) {
    int64_t v;
    return go_atoi(s, &v) ? v : 0;
}
/* filter.go:84-86 Uint64ToInt64: decimal string round trip; >= 2^63 -> ErrRange -> 0 */
static int64_t uint64_to_int64(uint64_t u) {
    char buf[32];
    snprintf(buf, sizeof buf, "%llu", (unsigned long long)u);
    return str_to_int64(buf);
}

uint64_t ora_str_to_uint64(const char *s) { return str_to_uint64(s); }
int64_t ora_uint64_to_int64(uint64_t u) { return uint64_to_int64(u); }

/* ------------------------------------------------------------------------------------------
 * Data contract ([3P] SCV api/v1 types as forced by the call sites)
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    char health[16]; /* Card.Health string, compared with "Healthy" (filter.go:53,57) */
    uint64_t free_memory, total_memory;      /* uint64 */
    uint64_t clock, core, bandwidth, power;  /* Go uint */
} ora_card;

typedef struct {
    char *key, *val;
} ora_kv;

typedef struct {
    uint32_t n;
    ora_kv *kv;
} ora_labels; /* map[string]string; a missing key is "label absent" */

typedef struct {
    int has_scv;              /* cache.Get succeeded (scheduler.go:80-84) */
    uint64_t card_number;     /* Status.CardNumber */
    uint32_t n_cards;         /* len(Status.CardList) -- independent of card_number */
    ora_card *cards;
    uint64_t free_memory_sum, total_memory_sum;
    uint32_t n_pods;          /* NodeInfo.Pods (algorithm.go:76) */
    ora_labels *pods;
} ora_node;

typedef struct {
    uint32_t n_nodes;
    ora_node *nodes;
} ora_cluster;

typedef struct {
    uint64_t max_bandwidth, max_clock, max_core, max_free_memory, max_power, max_total_memory;
} ora_max; /* collection.go:14-21, same field order */

static const char *label_get(const ora_labels *l, const char *key) {
    for (uint32_t i = 0; i < l->n; ++i)
        if (strcmp(l->kv[i].key, key) == 0) return l->kv[i].val;
    return NULL;
}

static void labels_free(ora_labels *l) {
    for (uint32_t i = 0; i < l->n; ++i) {
        free(l->kv[i].key);
        free(l->kv[i].val);
    }
    free(l->kv);
    l->kv = NULL;
    l->n = 0;
}

static void labels_set(ora_labels *l, uint32_t n, const char *const *keys, const char *const *vals) {
    labels_free(l);
    l->n = n;
    l->kv = n ? (ora_kv *)calloc(n, sizeof(ora_kv)) : NULL;
    for (uint32_t i = 0; i < n; ++i) {
        l->kv[i].key = strdup(keys[i]);
        l->kv[i].val = strdup(vals[i]);
    }
}

/* ------------------------------------------------------------------------------------------
 * filter.go
 * ---------------------------------------------------------------------------------------- */
/* filter.go:52-54 */
static int card_fits_memory(uint64_t memory, const ora_card *c) {
    return strcmp(c->health, "Healthy") == 0 && c->free_memory >= memory;
}
/* filter.go:56-58 (equality, not >=) */
static int card_fits_clock(uint64_t clock, const ora_card *c) {
    return strcmp(c->health, "Healthy") == 0 && c->clock == clock;
}
/* filter.go:11-16 */
static int pod_fits_number(const ora_labels *pod, const ora_node *s, uint64_t *number) {
    const char *v = label_get(pod, "scv/number");
    if (v) {
        *number = str_to_uint64(v);
        return str_to_uint64(v) <= s->card_number;
    }
    *number = 1;
    return s->card_number > 0;
}
/* filter.go:18-33 */
static int pod_fits_memory(uint64_t number, const ora_labels *pod, const ora_node *s, uint64_t *memory) {
    const char *v = label_get(pod, "scv/memory");
    if (v) {
        uint64_t fits = 0, m = str_to_uint64(v);
        for (uint32_t i = 0; i < s->n_cards; ++i)
            if (card_fits_memory(m, &s->cards[i])) fits++;
        *memory = m;
        return fits >= number;
    }
    *memory = 0;
    return 1;
}
/* filter.go:35-50 */
static int pod_fits_clock(uint64_t number, const ora_labels *pod, const ora_node *s, uint64_t *clock) {
    const char *v = label_get(pod, "scv/clock");
    if (v) {
        uint64_t fits = 0, c = str_to_uint64(v);
        for (uint32_t i = 0; i < s->n_cards; ++i)
            if (card_fits_clock(c, &s->cards[i])) fits++;
        *clock = c;
        return fits >= number;
    }
    *clock = 0;
    return 1;
}

/* scheduler.go:76-93 Yoda.Filter: 1 = Success, 0 = Unschedulable (missing SCV :81-84) */
static int yoda_filter(const ora_labels *pod, const ora_node *node) {
    if (!node->has_scv) return 0;
    uint64_t number, tmp;
    if (pod_fits_number(pod, node, &number)) {
        int fm = pod_fits_memory(number, pod, node, &tmp);
        int fc = pod_fits_clock(number, pod, node, &tmp);
        if (fm && fc) return 1;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * collection.go
 * ---------------------------------------------------------------------------------------- */
/* collection.go:59-78 */
static void process_max_value_with_card(const ora_card *c, ora_max *d) {
    if (c->free_memory > d->max_free_memory) d->max_free_memory = c->free_memory;
    if (c->clock > d->max_clock) d->max_clock = c->clock;
    if (c->total_memory > d->max_total_memory) d->max_total_memory = c->total_memory;
    if (c->bandwidth > d->max_bandwidth) d->max_bandwidth = c->bandwidth;
    if (c->core > d->max_core) d->max_core = c->core;
    if (c->power > d->max_power) d->max_power = c->power;
}

/* An Scv deep copy as the reference pays twice per pod (scheduler.go:98, collection.go:40).
 * Only used by the "faithful" timing driver; results do not depend on it. */
static ora_node scv_deep_copy(const ora_node *s) {
    ora_node c = *s;
    c.cards = (ora_card *)malloc(sizeof(ora_card) * (s->n_cards ? s->n_cards : 1));
    memcpy(c.cards, s->cards, sizeof(ora_card) * s->n_cards);
    return c;
}

/* collection.go:30-57 over the ScvList (= every node that has an SCV object) */
static void collect_max_values(const ora_labels *pod, const ora_cluster *cl, int faithful, ora_max *out) {
    ora_max d = {1, 1, 1, 1, 1, 1}; /* :31-38 */
    for (uint32_t n = 0; n < cl->n_nodes; ++n) {
        const ora_node *item = &cl->nodes[n];
        if (!item->has_scv) continue; /* not in the cache's list */
        ora_node copy;
        const ora_node *s = item;
        if (faithful) {
            copy = scv_deep_copy(item); /* :40 */
            s = &copy;
        }
        uint64_t number, memory, clock;
        if (pod_fits_number(pod, s, &number)) {
            int fm = pod_fits_memory(number, pod, s, &memory);
            int fc = pod_fits_clock(number, pod, s, &clock);
            if (fc && fm) {
                for (uint32_t i = 0; i < s->n_cards; ++i) {
                    const ora_card *c = &s->cards[i];
                    if (c->free_memory >= memory && c->clock >= clock) /* :46, no health test */
                        process_max_value_with_card(c, &d);
                }
            }
        }
        if (faithful) free(copy.cards);
    }
    *out = d;
}

/* ------------------------------------------------------------------------------------------
 * algorithm.go
 * ---------------------------------------------------------------------------------------- */
enum { BandwidthWeight = 1, ClockWeight = 1, CoreWeight = 1, PowerWeight = 1, FreeMemoryWeight = 2,
       TotalMemoryWeight = 1, ActualWeight = 2, AllocateWeight = 3 }; /* algorithm.go:16-26 */

/* algorithm.go:57-68; note :60 divides Clock by MaxBandwidth (sic) */
static uint64_t calculate_card_score(const ora_max *v, const ora_card *c) {
    uint64_t bandwidth = c->bandwidth * 100 / v->max_bandwidth;
    uint64_t clock = c->clock * 100 / v->max_bandwidth;
    uint64_t core = c->core * 100 / v->max_core;
    uint64_t power = c->power * 100 / v->max_power;
    uint64_t free_memory = c->free_memory * 100 / v->max_free_memory;
    uint64_t total_memory = c->total_memory * 100 / v->max_total_memory;
    return (bandwidth * BandwidthWeight + clock * ClockWeight + core * CoreWeight + power * PowerWeight) +
           free_memory * FreeMemoryWeight + total_memory * TotalMemoryWeight;
}

/* algorithm.go:41-55 */
static uint64_t calculate_basic_score(const ora_max *v, const ora_node *s, const ora_labels *pod) {
    uint64_t card_score = 0, number, memory, clock;
    if (pod_fits_number(pod, s, &number)) {
        int fm = pod_fits_memory(number, pod, s, &memory);
        int fc = pod_fits_clock(number, pod, s, &clock);
        if (fc && fm) {
            for (uint32_t i = 0; i < s->n_cards; ++i) {
                const ora_card *c = &s->cards[i];
                if (c->free_memory >= memory && c->clock >= clock) card_score += calculate_card_score(v, c);
            }
        }
    }
    return card_score;
}

/* algorithm.go:70-72; TotalMemorySum == 0 panics in Go */
static int calculate_actual_score(const ora_node *s, uint64_t *out) {
    if (s->total_memory_sum == 0) return ORA_ERR_DIV0;
    *out = (s->free_memory_sum * 100 / s->total_memory_sum) * ActualWeight;
    return ORA_OK;
}

/* algorithm.go:74-87 */
static int calculate_allocate_score(const ora_node *s, uint64_t *out) {
    uint64_t allocate_memory_sum = 0;
    for (uint32_t i = 0; i < s->n_pods; ++i) {
        const char *mem = label_get(&s->pods[i], "scv/memory");
        if (mem) allocate_memory_sum += str_to_uint64(mem);
    }
    if (s->total_memory_sum < allocate_memory_sum) {
        *out = 0;
        return ORA_OK;
    }
    if (s->total_memory_sum == 0) return ORA_ERR_DIV0;
    *out = (s->total_memory_sum - allocate_memory_sum) * 100 / s->total_memory_sum * AllocateWeight;
    return ORA_OK;
}

/* algorithm.go:28-39 CalculateScore + scheduler.go:128 Uint64ToInt64 (Yoda.Score :109-130) */
static int yoda_score(const ora_max *v, const ora_labels *pod, const ora_node *s, int64_t *out) {
    uint64_t alloc, actual;
    uint64_t basic = calculate_basic_score(v, s, pod);
    int rc = calculate_allocate_score(s, &alloc);
    if (rc) return rc;
    rc = calculate_actual_score(s, &actual);
    if (rc) return rc;
    *out = uint64_to_int64(basic + alloc + actual);
    return ORA_OK;
}

/* scheduler.go:132-157 Yoda.NormalizeScore, in place.  int64 arithmetic wraps in Go, so it is
 * done unsigned here and cast back; Go's MinInt64 / -1 == MinInt64 rule is kept. */
static int64_t go_div_i64(int64_t a, int64_t b) {
    if (a == INT64_MIN && b == -1) return a;
    return a / b;
}
int ora_normalize_score(int64_t *scores, uint32_t n) {
    if (n == 0) return ORA_ERR_ARG; /* scores[0] would panic (:135) */
    int64_t highest = 0, lowest = scores[0];
    for (uint32_t i = 0; i < n; ++i) {
        if (scores[i] < lowest) lowest = scores[i];
        if (scores[i] > highest) highest = scores[i];
    }
    if (highest == lowest) lowest = (int64_t)((uint64_t)lowest - 1);
    int64_t den = (int64_t)((uint64_t)highest - (uint64_t)lowest);
    for (uint32_t i = 0; i < n; ++i) {
        int64_t num = (int64_t)(((uint64_t)scores[i] - (uint64_t)lowest) * (uint64_t)100); /* MaxNodeScore */
        scores[i] = go_div_i64(num, den);
    }
    return ORA_OK;
}

/* ------------------------------------------------------------------------------------------
 * Cluster construction (test/bench plumbing, no reference counterpart)
 * ---------------------------------------------------------------------------------------- */
ora_cluster *ora_cluster_new(uint32_t n_nodes) {
    ora_cluster *c = (ora_cluster *)calloc(1, sizeof *c);
    c->n_nodes = n_nodes;
    c->nodes = (ora_node *)calloc(n_nodes ? n_nodes : 1, sizeof(ora_node));
    return c;
}

static void node_clear(ora_node *nd) {
    free(nd->cards);
    for (uint32_t i = 0; i < nd->n_pods; ++i) labels_free(&nd->pods[i]);
    free(nd->pods);
    memset(nd, 0, sizeof *nd);
}

void ora_cluster_free(ora_cluster *c) {
    if (!c) return;
    for (uint32_t i = 0; i < c->n_nodes; ++i) node_clear(&c->nodes[i]);
    free(c->nodes);
    free(c);
}

/* health[i] is the literal Health string of card i */
int ora_cluster_set_node(ora_cluster *c, uint32_t idx, int has_scv, uint64_t card_number, uint32_t n_cards,
                         const char *const *health, const uint64_t *free_memory, const uint64_t *total_memory,
                         const uint64_t *clock, const uint64_t *core, const uint64_t *bandwidth,
                         const uint64_t *power, uint64_t free_sum, uint64_t total_sum) {
    if (idx >= c->n_nodes) return ORA_ERR_ARG;
    ora_node *nd = &c->nodes[idx];
    free(nd->cards);
    nd->has_scv = has_scv;
    nd->card_number = card_number;
    nd->n_cards = n_cards;
    nd->cards = (ora_card *)calloc(n_cards ? n_cards : 1, sizeof(ora_card));
    for (uint32_t i = 0; i < n_cards; ++i) {
        ora_card *k = &nd->cards[i];
        strncpy(k->health, health[i], sizeof k->health - 1);
        k->free_memory = free_memory[i];
        k->total_memory = total_memory[i];
        k->clock = clock[i];
        k->core = core[i];
        k->bandwidth = bandwidth[i];
        k->power = power[i];
    }
    nd->free_memory_sum = free_sum;
    nd->total_memory_sum = total_sum;
    return ORA_OK;
}

/* Running pods on the node: pod i has n_kv[i] labels taken consecutively from keys/vals. */
int ora_cluster_set_node_pods(ora_cluster *c, uint32_t idx, uint32_t n_pods, const uint32_t *n_kv,
                              const char *const *keys, const char *const *vals) {
    if (idx >= c->n_nodes) return ORA_ERR_ARG;
    ora_node *nd = &c->nodes[idx];
    for (uint32_t i = 0; i < nd->n_pods; ++i) labels_free(&nd->pods[i]);
    free(nd->pods);
    nd->n_pods = n_pods;
    nd->pods = (ora_labels *)calloc(n_pods ? n_pods : 1, sizeof(ora_labels));
    uint32_t off = 0;
    for (uint32_t i = 0; i < n_pods; ++i) {
        labels_set(&nd->pods[i], n_kv[i], keys + off, vals + off);
        off += n_kv[i];
    }
    return ORA_OK;
}

/* Bulk fill from flat arrays ([n][8] card planes, healthy bitmask).  Running pods are given
 * as a CSR list of StrToUint64(scv/memory) values and rendered back to decimal label strings. */
int ora_cluster_fill(ora_cluster *c, const uint8_t *has_scv, const uint64_t *card_number, const uint8_t *n_cards,
                     const uint8_t *healthy_mask, const uint64_t *free_memory, const uint64_t *total_memory,
                     const uint64_t *clock, const uint64_t *core, const uint64_t *bandwidth, const uint64_t *power,
                     const uint64_t *free_sum, const uint64_t *total_sum, const uint32_t *pod_off,
                     const uint64_t *pod_mem) {
    static const char *H = "Healthy", *U = "Unhealthy";
    for (uint32_t n = 0; n < c->n_nodes; ++n) {
        const char *hs[8];
        if (n_cards[n] > 8) return ORA_ERR_ARG;
        for (int g = 0; g < 8; ++g) hs[g] = ((healthy_mask[n] >> g) & 1) ? H : U;
        ora_cluster_set_node(c, n, has_scv[n], card_number[n], n_cards[n], hs, free_memory + 8 * (size_t)n,
                             total_memory + 8 * (size_t)n, clock + 8 * (size_t)n, core + 8 * (size_t)n,
                             bandwidth + 8 * (size_t)n, power + 8 * (size_t)n, free_sum[n], total_sum[n]);
        uint32_t np = pod_off ? pod_off[n + 1] - pod_off[n] : 0;
        if (np) {
            uint32_t *nkv = (uint32_t *)malloc(sizeof(uint32_t) * np);
            const char **keys = (const char **)malloc(sizeof(char *) * np);
            char **vals = (char **)malloc(sizeof(char *) * np);
            for (uint32_t i = 0; i < np; ++i) {
                nkv[i] = 1;
                keys[i] = "scv/memory";
                vals[i] = (char *)malloc(24);
                /* signed rendering, so that values >= 2^63 (labels such as "-5") survive Atoi */
                snprintf(vals[i], 24, "%lld", (long long)pod_mem[pod_off[n] + i]);
            }
            ora_cluster_set_node_pods(c, n, np, nkv, keys, (const char *const *)vals);
            for (uint32_t i = 0; i < np; ++i) free(vals[i]);
            free(nkv);
            free(keys);
            free(vals);
        } else {
            ora_cluster_set_node_pods(c, n, 0, NULL, NULL, NULL);
        }
    }
    return ORA_OK;
}

/* ------------------------------------------------------------------------------------------
 * Batch driver: the scheduling cycle of SURVEY.md section 3 ("intended order"), for each pod on
 * the same frozen snapshot:
 *   Filter(all nodes) -> CollectMaxValues(all SCVs) -> Score(feasible) -> NormalizeScore
 *   -> chosen = lowest index among nodes with the maximum normalized score.
 * Outputs (any pointer may be NULL):
 *   feasible [P][N] u8; raw [P][N] i64 = Score() for every node that has an SCV (0 otherwise);
 *   norm [P][N] i64 (0 for infeasible); max6 [P][6] in collection.go:14-21 order;
 *   lowest/highest [P] as used at scheduler.go:152 (after the lowest-- adjustment);
 *   count [P]; chosen [P] (-1 when nothing is feasible).
 * faithful != 0 additionally performs the per-call Scv copies the Go plugin pays
 * (scheduler.go:80,98,118; collection.go:40) -- same results, used for the timed baseline.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    const ora_cluster *cl;
    uint32_t n_pods;
    const ora_labels *pods;
    int faithful;
    uint8_t *feasible;
    int64_t *raw, *norm;
    uint64_t *max6;
    int64_t *lowest, *highest;
    uint32_t *count;
    int32_t *chosen;
    int rc;
    uint32_t next; /* shared work counter */
    pthread_mutex_t mu;
} ora_job;

static int eval_one_pod(ora_job *j, uint32_t p, uint8_t *feas, int64_t *raw, int64_t *list) {
    const ora_cluster *cl = j->cl;
    const ora_labels *pod = &j->pods[p];
    const uint32_t N = cl->n_nodes;
    uint32_t cnt = 0;
    /* Filter, once per node (scheduler.go:76-93) */
    for (uint32_t n = 0; n < N; ++n) {
        const ora_node *nd = &cl->nodes[n];
        if (j->faithful && nd->has_scv) {
            ora_node copy = scv_deep_copy(nd); /* cache.Get deep copy, :80 */
            feas[n] = (uint8_t)yoda_filter(pod, &copy);
            free(copy.cards);
        } else {
            feas[n] = (uint8_t)yoda_filter(pod, nd);
        }
        cnt += feas[n];
    }
    /* PostFilter/CollectMaxValues (scheduler.go:95-103) */
    ora_max mx;
    collect_max_values(pod, cl, j->faithful, &mx);
    /* Score every node that has an SCV (scheduler.go:109-130) */
    for (uint32_t n = 0; n < N; ++n) {
        const ora_node *nd = &cl->nodes[n];
        raw[n] = 0;
        if (!nd->has_scv) continue;
        int rc;
        if (j->faithful) {
            ora_node copy = scv_deep_copy(nd); /* :117-118 */
            copy.pods = nd->pods;
            rc = yoda_score(&mx, pod, &copy, &raw[n]);
            free(copy.cards);
        } else {
            rc = yoda_score(&mx, pod, nd, &raw[n]);
        }
        if (rc) return rc;
    }
    /* NormalizeScore over the feasible list (scheduler.go:132-157) */
    uint32_t k = 0;
    for (uint32_t n = 0; n < N; ++n)
        if (feas[n]) list[k++] = raw[n];
    int64_t hi = 0, lo = 0;
    if (k) {
        lo = list[0];
        for (uint32_t i = 0; i < k; ++i) {
            if (list[i] < lo) lo = list[i];
            if (list[i] > hi) hi = list[i];
        }
        if (hi == lo) lo = (int64_t)((uint64_t)lo - 1);
        ora_normalize_score(list, k);
    }
    int32_t chosen = -1;
    int64_t best = -1;
    k = 0;
    for (uint32_t n = 0; n < N; ++n) {
        int64_t nv = 0;
        if (feas[n]) {
            nv = list[k++];
            if (nv > best) {
                best = nv;
                chosen = (int32_t)n;
            }
        }
        if (j->norm) j->norm[(size_t)p * N + n] = nv;
    }
    if (j->feasible) memcpy(j->feasible + (size_t)p * N, feas, N);
    if (j->raw) memcpy(j->raw + (size_t)p * N, raw, sizeof(int64_t) * N);
    if (j->max6) {
        uint64_t *m = j->max6 + (size_t)p * 6;
        m[0] = mx.max_bandwidth; m[1] = mx.max_clock; m[2] = mx.max_core;
        m[3] = mx.max_free_memory; m[4] = mx.max_power; m[5] = mx.max_total_memory;
    }
    if (j->lowest) j->lowest[p] = lo;
    if (j->highest) j->highest[p] = hi;
    if (j->count) j->count[p] = cnt;
    if (j->chosen) j->chosen[p] = chosen;
    return ORA_OK;
}

static void *worker(void *arg) {
    ora_job *j = (ora_job *)arg;
    const uint32_t N = j->cl->n_nodes;
    uint8_t *feas = (uint8_t *)malloc(N ? N : 1);
    int64_t *raw = (int64_t *)malloc(sizeof(int64_t) * (N ? N : 1));
    int64_t *list = (int64_t *)malloc(sizeof(int64_t) * (N ? N : 1));
    for (;;) {
        pthread_mutex_lock(&j->mu);
        uint32_t p = j->next++;
        pthread_mutex_unlock(&j->mu);
        if (p >= j->n_pods) break;
        int rc = eval_one_pod(j, p, feas, raw, list);
        if (rc) {
            pthread_mutex_lock(&j->mu);
            j->rc = rc;
            pthread_mutex_unlock(&j->mu);
        }
    }
    free(feas);
    free(raw);
    free(list);
    return NULL;
}

/* Pod label values: NULL = label absent. */
int ora_eval_batch(const ora_cluster *cl, uint32_t n_pods, const char *const *number, const char *const *memory,
                   const char *const *clock, int faithful, int threads, uint8_t *feasible, int64_t *raw,
                   int64_t *norm, uint64_t *max6, int64_t *lowest, int64_t *highest, uint32_t *count,
                   int32_t *chosen) {
    ora_labels *pods = (ora_labels *)calloc(n_pods ? n_pods : 1, sizeof(ora_labels));
    for (uint32_t p = 0; p < n_pods; ++p) {
        const char *keys[3], *vals[3];
        uint32_t k = 0;
        if (memory && memory[p]) { keys[k] = "scv/memory"; vals[k++] = memory[p]; }
        if (number && number[p]) { keys[k] = "scv/number"; vals[k++] = number[p]; }
        if (clock && clock[p]) { keys[k] = "scv/clock"; vals[k++] = clock[p]; }
        labels_set(&pods[p], k, keys, vals);
    }
    ora_job j;
    memset(&j, 0, sizeof j);
    j.cl = cl; j.n_pods = n_pods; j.pods = pods; j.faithful = faithful;
    j.feasible = feasible; j.raw = raw; j.norm = norm; j.max6 = max6;
    j.lowest = lowest; j.highest = highest; j.count = count; j.chosen = chosen;
    pthread_mutex_init(&j.mu, NULL);
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    if (threads == 1) {
        worker(&j);
    } else {
        pthread_t th[256];
        for (int t = 0; t < threads; ++t) pthread_create(&th[t], NULL, worker, &j);
        for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
    }
    pthread_mutex_destroy(&j.mu);
    for (uint32_t p = 0; p < n_pods; ++p) labels_free(&pods[p]);
    free(pods);
    return j.rc;
}

/* Single-call entry points for unit KATs ------------------------------------------------ */
int ora_filter(const ora_cluster *cl, uint32_t node, const char *number, const char *memory, const char *clock) {
    const char *keys[3], *vals[3];
    uint32_t k = 0;
    ora_labels pod = {0, NULL};
    if (memory) { keys[k] = "scv/memory"; vals[k++] = memory; }
    if (number) { keys[k] = "scv/number"; vals[k++] = number; }
    if (clock) { keys[k] = "scv/clock"; vals[k++] = clock; }
    labels_set(&pod, k, keys, vals);
    int r = yoda_filter(&pod, &cl->nodes[node]);
    labels_free(&pod);
    return r;
}

/* static (pod-independent) score parts of a node: allocate (algorithm.go:74-87) and actual (:70-72) */
int ora_node_static(const ora_cluster *cl, uint32_t node, uint64_t *alloc, uint64_t *actual) {
    int rc = calculate_allocate_score(&cl->nodes[node], alloc);
    if (rc) return rc;
    return calculate_actual_score(&cl->nodes[node], actual);
}

/* Score of every node of `cl` for one pod under GIVEN maxima (a node shard scoring with the
 * cluster-wide collection.MaxValue).  raw[n] as in ora_eval_batch; feasible[n] = Filter verdict. */
int ora_score_with_max(const ora_cluster *cl, const char *number, const char *memory, const char *clock,
                       const uint64_t *max6, uint8_t *feasible, int64_t *raw) {
    const char *keys[3], *vals[3];
    uint32_t k = 0;
    ora_labels pod = {0, NULL};
    if (memory) { keys[k] = "scv/memory"; vals[k++] = memory; }
    if (number) { keys[k] = "scv/number"; vals[k++] = number; }
    if (clock) { keys[k] = "scv/clock"; vals[k++] = clock; }
    labels_set(&pod, k, keys, vals);
    ora_max mx = {max6[0], max6[1], max6[2], max6[3], max6[4], max6[5]};
    int rc = ORA_OK;
    for (uint32_t n = 0; n < cl->n_nodes && !rc; ++n) {
        const ora_node *nd = &cl->nodes[n];
        feasible[n] = (uint8_t)yoda_filter(&pod, nd);
        raw[n] = 0;
        if (nd->has_scv) rc = yoda_score(&mx, &pod, nd, &raw[n]);
    }
    labels_free(&pod);
    return rc;
}

/* sort.go:8-18: Less(a, b) = priority(a) > priority(b), priority = Atoi(label scv/priority) with the
 * error dropped (0 when the label is absent).  order[] = stable insertion order under Less. */
static int64_t pod_priority(const char *label) {
    if (!label) return 0;
    /* strconv.Atoi returns the clamped value together with a range error; GetPodPriority keeps the value */
    size_t n = strlen(label), i = 0;
    int neg = 0;
    if (n == 0) return 0;
    if (label[0] == '+' || label[0] == '-') { neg = label[0] == '-'; i = 1; if (n == 1) return 0; }
    unsigned __int128 acc = 0;
    int over = 0;
    for (; i < n; ++i) {
        unsigned d = (unsigned char)label[i] - (unsigned)'0';
        if (d > 9) return 0; /* syntax error -> 0 */
        if (!over) { acc = acc * 10 + d; if (acc > ((unsigned __int128)1 << 63)) over = 1; }
    }
    if (!neg && acc > (((unsigned __int128)1 << 63) - 1)) over = 1;
    if (over) return neg ? INT64_MIN : INT64_MAX;
    return neg ? (int64_t)(0 - (uint64_t)acc) : (int64_t)(uint64_t)acc;
}
int64_t ora_pod_priority(const char *label) { return pod_priority(label); }
void ora_queue_order(uint32_t n, const char *const *priority_labels, uint32_t *order) {
    int64_t *pri = (int64_t *)malloc(sizeof(int64_t) * (n ? n : 1));
    for (uint32_t i = 0; i < n; ++i) pri[i] = pod_priority(priority_labels[i]);
    for (uint32_t i = 0; i < n; ++i) { /* insertion sort: move i before every earlier pod it is Less-than */
        uint32_t k = i;
        while (k > 0 && pri[i] > pri[order[k - 1]]) { order[k] = order[k - 1]; --k; }
        order[k] = i;
    }
    free(pri);
}

/* Sequential-commit batch: a run of single-pod scheduling cycles.  Pods are taken in `order` (NULL =
 * as given); after each cycle the pod is "assumed" onto its chosen node, i.e. appended to that node's
 * NodeInfo.Pods, which is what CalculateAllocateScore (algorithm.go:74-87) reads in the next cycle.
 * The cluster is restored before returning.  Outputs indexed by pod (not by order). */
int ora_eval_sequential(ora_cluster *cl, uint32_t n_pods, const char *const *number, const char *const *memory,
                        const char *const *clock, const uint32_t *order, uint64_t *max6, int64_t *lowest, int64_t *highest,
                        uint32_t *count, int32_t *chosen) {
    const uint32_t N = cl->n_nodes;
    uint32_t *orig_np = (uint32_t *)malloc(sizeof(uint32_t) * (N ? N : 1));
    for (uint32_t n = 0; n < N; ++n) orig_np[n] = cl->nodes[n].n_pods;
    ora_labels *pods = (ora_labels *)calloc(n_pods ? n_pods : 1, sizeof(ora_labels));
    for (uint32_t p = 0; p < n_pods; ++p) {
        const char *keys[3], *vals[3];
        uint32_t k = 0;
        if (memory && memory[p]) { keys[k] = "scv/memory"; vals[k++] = memory[p]; }
        if (number && number[p]) { keys[k] = "scv/number"; vals[k++] = number[p]; }
        if (clock && clock[p]) { keys[k] = "scv/clock"; vals[k++] = clock[p]; }
        labels_set(&pods[p], k, keys, vals);
    }
    ora_job j;
    memset(&j, 0, sizeof j);
    j.cl = cl; j.n_pods = n_pods; j.pods = pods;
    j.max6 = max6; j.lowest = lowest; j.highest = highest; j.count = count; j.chosen = chosen;
    uint8_t *feas = (uint8_t *)malloc(N ? N : 1);
    int64_t *raw = (int64_t *)malloc(sizeof(int64_t) * (N ? N : 1));
    int64_t *list = (int64_t *)malloc(sizeof(int64_t) * (N ? N : 1));
    int rc = ORA_OK;
    for (uint32_t k = 0; k < n_pods && !rc; ++k) {
        const uint32_t p = order ? order[k] : k;
        rc = eval_one_pod(&j, p, feas, raw, list);
        if (!rc && chosen[p] >= 0) { /* assume: the pod now runs on the chosen node */
            ora_node *nd = &cl->nodes[chosen[p]];
            nd->pods = (ora_labels *)realloc(nd->pods, sizeof(ora_labels) * (nd->n_pods + 1));
            memset(&nd->pods[nd->n_pods], 0, sizeof(ora_labels));
            const char *keys[3], *vals[3];
            for (uint32_t q = 0; q < pods[p].n; ++q) { keys[q] = pods[p].kv[q].key; vals[q] = pods[p].kv[q].val; }
            labels_set(&nd->pods[nd->n_pods], pods[p].n, keys, vals);
            nd->n_pods++;
        }
    }
    for (uint32_t n = 0; n < N; ++n) { /* restore */
        ora_node *nd = &cl->nodes[n];
        while (nd->n_pods > orig_np[n]) labels_free(&nd->pods[--nd->n_pods]);
    }
    for (uint32_t p = 0; p < n_pods; ++p) labels_free(&pods[p]);
    free(pods); free(feas); free(raw); free(list); free(orig_np);
    return rc;
}

/* ------------------------------------------------------------------------------------------------
 * Framework tail: what kube-scheduler v1.20.0 does around the plugin for one pod.
 * [3P] k8s.io/kubernetes v1.20.0 pkg/scheduler/core/generic_scheduler.go is not under /root/reference
 * (un-vendored module pinned at go.mod:12); restated from its published algorithm:
 *   numFeasibleNodesToFind: all nodes if < 100 of them or percentage >= 100; percentage <= 0 -> adaptive
 *     50 - numAllNodes/125, floor 5; result numAllNodes*percentage/100, floor 100.
 *   findNodesThatPassFilters: nodes are visited round-robin from nextStartNodeIndex until that many feasible
 *     ones are found; nextStartNodeIndex advances by the number of nodes visited.  (Upstream visits them from
 *     16 goroutines, so which nodes make the cut is a race; this restatement -- like the device version --
 *     takes them strictly in visiting order, which is one of the outcomes the race can produce.)
 *   prioritizeNodes: Score + NormalizeScore over the nodes found (scheduler.go:109-157 for Yoda), times the
 *     plugin weight (300, deploy/yoda-scheduler.yaml:30), plus the other plugins' weighted scores (here: an
 *     optional per-node addend).  With exactly one feasible node scoring is skipped.
 *   selectHost: highest total, ties broken by reservoir sampling -- rand.Intn(cnt) == 0 replaces the
 *     candidate; here rand.Intn(k) is SplitMix64(seed, pod, k) % k so that runs can be compared.
 * CollectMaxValues still sees every SCV of the cluster (collection.go:39-51): sampling does not touch it.
 * ------------------------------------------------------------------------------------------------ */
static uint64_t tail_sm64(uint64_t x) {
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint64_t tail_rand(uint64_t seed, uint64_t pod, uint64_t k) { return tail_sm64(tail_sm64(tail_sm64(seed) + pod) + k); }

uint32_t ora_num_feasible_nodes_to_find(uint32_t num_all_nodes, int32_t percentage) {
    const int32_t all = (int32_t)num_all_nodes;
    if (all < 100 || percentage >= 100) return num_all_nodes;
    int32_t adaptive = percentage;
    if (adaptive <= 0) {
        adaptive = 50 - all / 125;
        if (adaptive < 5) adaptive = 5;
    }
    const int32_t num = (int32_t)((int64_t)all * adaptive / 100);
    return num < 100 ? 100u : (uint32_t)num;
}

/* out arrays are indexed by pod: chosen (-1 = no feasible node), sampled (feasible nodes found), processed (nodes
 * visited), start (the pod's own start index), lowest/highest (of the sampled list, after lowest--; 0/0 when scoring
 * was skipped), top (winning total), ties (nodes holding it).  chain != 0: pods are consecutive cycles, each starting
 * where the previous stopped; otherwise every pod starts at start_index.  other: per-node addend or NULL. */
int ora_framework_tail(const ora_cluster *cl, uint32_t n_pods, const char *const *number, const char *const *memory,
                       const char *const *clock, uint32_t weight, int32_t percentage, uint64_t seed, uint32_t start_index,
                       int chain, const int64_t *other, int32_t *chosen, uint32_t *sampled, uint32_t *processed,
                       uint32_t *start, int64_t *lowest, int64_t *highest, int64_t *top, uint32_t *ties) {
    const uint32_t N = cl->n_nodes;
    const uint32_t want = ora_num_feasible_nodes_to_find(N, percentage);
    uint32_t *found = (uint32_t *)malloc(sizeof(uint32_t) * (N ? N : 1));
    int64_t *list = (int64_t *)malloc(sizeof(int64_t) * (N ? N : 1));
    uint32_t next_start = N ? start_index % N : 0;
    int rc = ORA_OK;
    for (uint32_t p = 0; p < n_pods && !rc; ++p) {
        ora_labels pod;
        memset(&pod, 0, sizeof pod);
        const char *keys[3], *vals[3];
        uint32_t k = 0;
        if (memory && memory[p]) { keys[k] = "scv/memory"; vals[k++] = memory[p]; }
        if (number && number[p]) { keys[k] = "scv/number"; vals[k++] = number[p]; }
        if (clock && clock[p]) { keys[k] = "scv/clock"; vals[k++] = clock[p]; }
        labels_set(&pod, k, keys, vals);
        const uint32_t s0 = chain ? next_start : (N ? start_index % N : 0);
        uint32_t nf = 0, visited = 0;
        for (uint32_t i = 0; i < N && nf < want; ++i) { /* findNodesThatPassFilters */
            const uint32_t n = (s0 + i) % N;
            visited++;
            if (yoda_filter(&pod, &cl->nodes[n])) found[nf++] = n;
        }
        next_start = N ? (s0 + visited) % N : 0;
        start[p] = s0; sampled[p] = nf; processed[p] = visited;
        lowest[p] = highest[p] = 0; top[p] = 0; ties[p] = 0; chosen[p] = -1;
        if (nf == 1) { /* a single feasible node is taken without scoring */
            chosen[p] = (int32_t)found[0];
            ties[p] = 1;
        } else if (nf > 1) {
            ora_max mx;
            collect_max_values(&pod, cl, 0, &mx);
            for (uint32_t i = 0; i < nf && !rc; ++i) rc = yoda_score(&mx, &pod, &cl->nodes[found[i]], &list[i]);
            if (!rc) {
                int64_t hi = 0, lo = list[0];
                for (uint32_t i = 0; i < nf; ++i) {
                    if (list[i] < lo) lo = list[i];
                    if (list[i] > hi) hi = list[i];
                }
                if (hi == lo) lo = (int64_t)((uint64_t)lo - 1);
                lowest[p] = lo; highest[p] = hi;
                ora_normalize_score(list, nf);
                /* weight, other plugins, selectHost */
                int64_t max_score = 0;
                uint32_t cnt = 0;
                for (uint32_t i = 0; i < nf; ++i) {
                    const int64_t total = (int64_t)((uint64_t)list[i] * (uint64_t)weight + (uint64_t)(other ? other[found[i]] : 0));
                    if (i == 0 || total > max_score) {
                        max_score = total;
                        chosen[p] = (int32_t)found[i];
                        cnt = 1;
                    } else if (total == max_score) {
                        cnt++;
                        if (tail_rand(seed, p, cnt) % cnt == 0) chosen[p] = (int32_t)found[i];
                    }
                }
                top[p] = max_score; ties[p] = cnt;
            }
        }
        labels_free(&pod);
    }
    free(found); free(list);
    return rc;
}
