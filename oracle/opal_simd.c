/*
 * oracle/opal_simd.c -- TEST / BENCH INFRASTRUCTURE ONLY: a CPU baseline with
 * the parallelisation scheme of the reference's engine.
 *
 * The reference's DP engine (Opal, absent from /root/reference: empty
 * vendor/opal submodule) is an inter-sequence SIMD Smith-Waterman in the style
 * of SWIPE: every SIMD lane holds a different database sequence
 * (/root/reference/README.md:26-28), scores are kept in narrow integers and a
 * sequence whose score overflows is recomputed at a wider type
 * (lib.pyx:1283-1289).  This file restates that scheme in portable C -- LANES
 * 16-bit lanes over LANES targets of similar length, written as plain lane
 * loops that gcc -O3 -march=native turns into AVX2 / AVX-512 -- so that
 * bench.py's CPU arm is not a scalar strawman.  Score mode, local alignment
 * only; overflowing targets are recomputed by the scalar oracle
 * (opal_oracle.c).  Results are checked against the scalar oracle in
 * tests/test_oracle_golden.py.
 */
#include "opal_oracle.h"

#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define LANES 32
#define PADSCORE (-64)

typedef struct {
    const unsigned char* q;
    int m;
    const unsigned char* residues;
    const long long* offsets;
    const int* lengths;
    const int* order; /* target indices sorted by length, descending */
    int n;
    int go, ge;
    const int* S;
    int A;
    int* scores;
    int* next_batch;
    int nbatches;
    pthread_mutex_t* mu;
} Job;

/* one batch of up to LANES targets */
static void batch_sw(const Job* J, int b, int16_t* restrict H, int16_t* restrict E, int16_t* restrict prof) {
    const int m = J->m, A = J->A, go = J->go, ge = J->ge;
    const int first = b * LANES;
    const int cnt = (J->n - first < LANES) ? (J->n - first) : LANES;
    const unsigned char* seq[LANES];
    int len[LANES];
    int maxlen = 0;
    for (int l = 0; l < LANES; l++) {
        if (l < cnt) {
            const int t = J->order[first + l];
            seq[l] = J->residues + J->offsets[t];
            len[l] = J->lengths[t];
        } else {
            seq[l] = NULL;
            len[l] = 0;
        }
        if (len[l] > maxlen) maxlen = len[l];
    }
    int smax = 0;
    for (int i = 0; i < A * A; i++) if (J->S[i] > smax) smax = J->S[i];
    const int limit = 32767 - smax - 1;
    memset(H, 0, sizeof(int16_t) * (size_t)m * LANES);
    for (size_t x = 0; x < (size_t)m * LANES; x++) E[x] = -16384;
    int16_t best[LANES];
    for (int l = 0; l < LANES; l++) best[l] = 0;

    for (int j = 0; j < maxlen; j++) {
        /* score of every alphabet letter against this column's residue of every lane */
        for (int a = 0; a < A; a++) {
            int16_t* restrict p = prof + (size_t)a * LANES;
            for (int l = 0; l < LANES; l++) p[l] = (j < len[l]) ? (int16_t)J->S[a * A + seq[l][j]] : (int16_t)PADSCORE;
        }
        int16_t f[LANES], hup[LANES], hdiag[LANES];
        for (int l = 0; l < LANES; l++) { f[l] = -16384; hup[l] = 0; hdiag[l] = 0; }
        for (int i = 0; i < m; i++) {
            int16_t* restrict Hi = H + (size_t)i * LANES;
            int16_t* restrict Ei = E + (size_t)i * LANES;
            const int16_t* restrict p = prof + (size_t)J->q[i] * LANES;
            for (int l = 0; l < LANES; l++) {
                const int16_t hl = Hi[l];
                int16_t e = (int16_t)(Ei[l] - ge), e2 = (int16_t)(hl - go);
                e = e > e2 ? e : e2;
                int16_t ff = (int16_t)(f[l] - ge), f2 = (int16_t)(hup[l] - go);
                ff = ff > f2 ? ff : f2;
                int16_t h = (int16_t)(hdiag[l] + p[l]);
                h = h > e ? h : e;
                h = h > ff ? h : ff;
                h = h > 0 ? h : 0;
                hdiag[l] = hl;
                Hi[l] = h;
                Ei[l] = e;
                f[l] = ff;
                hup[l] = h;
                best[l] = best[l] > h ? best[l] : h;
            }
        }
        /* keep E / F away from the int16 floor (they only ever decay by ext per step) */
        if ((j & 1023) == 1023)
            for (size_t x = 0; x < (size_t)m * LANES; x++) if (E[x] < -16384) E[x] = -16384;
    }
    for (int l = 0; l < cnt; l++) {
        const int t = J->order[first + l];
        if (best[l] >= limit) {
            /* overflow: recompute at full width with the scalar oracle */
            int one_len = J->lengths[t];
            long long one_off = 0;
            int sc = 0;
            oracle_search_flat(J->q, m, J->residues + J->offsets[t], &one_off, &one_len, 1, go, ge, J->S, A,
                               ORACLE_SEARCH_SCORE, ORACLE_MODE_SW, &sc, NULL, NULL, NULL, NULL, NULL, NULL, NULL, 1);
            J->scores[t] = sc;
        } else {
            J->scores[t] = best[l];
        }
    }
}

static void* worker(void* arg) {
    const Job* J = (const Job*)arg;
    int16_t* H = (int16_t*)aligned_alloc(64, sizeof(int16_t) * (size_t)(J->m > 0 ? J->m : 1) * LANES);
    int16_t* E = (int16_t*)aligned_alloc(64, sizeof(int16_t) * (size_t)(J->m > 0 ? J->m : 1) * LANES);
    int16_t* prof = (int16_t*)aligned_alloc(64, sizeof(int16_t) * (size_t)J->A * LANES);
    for (;;) {
        pthread_mutex_lock(J->mu);
        const int b = (*J->next_batch)++;
        pthread_mutex_unlock(J->mu);
        if (b >= J->nbatches) break;
        batch_sw(J, b, H, E, prof);
    }
    free(H); free(E); free(prof);
    return NULL;
}

static const int* g_len_for_sort;
static int by_length_desc(const void* a, const void* b) {
    const int x = *(const int*)a, y = *(const int*)b;
    const int lx = g_len_for_sort[x], ly = g_len_for_sort[y];
    if (lx != ly) return (lx < ly) - (lx > ly);
    return (x > y) - (x < y);
}

/* Smith-Waterman scores of one query against n packed targets. Returns 0. */
int oracle_sw_scores_simd(const unsigned char* query, int m, const unsigned char* residues, const long long* offsets,
                          const int* lengths, int n, int go, int ge, const int* S, int A, int* scores, int threads) {
    if (n <= 0) return 0;
    if (m <= 0) { for (int t = 0; t < n; t++) scores[t] = 0; return 0; }
    int* order = (int*)malloc(sizeof(int) * (size_t)n);
    for (int t = 0; t < n; t++) order[t] = t;
    g_len_for_sort = lengths;
    qsort(order, (size_t)n, sizeof(int), by_length_desc);
    pthread_mutex_t mu;
    pthread_mutex_init(&mu, NULL);
    int next = 0;
    Job J = {query, m, residues, offsets, lengths, order, n, go, ge, S, A, scores, &next, (n + LANES - 1) / LANES, &mu};
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    pthread_t th[256];
    for (int t = 0; t < threads; t++) pthread_create(&th[t], NULL, worker, &J);
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
    pthread_mutex_destroy(&mu);
    free(order);
    return 0;
}
