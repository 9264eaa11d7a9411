/*
 * wire_check.c -- pins the wire format of sb_iter_in without a JVM (built and run by tests/test_wire.py).
 *
 * It writes a DTMasterParams message byte by byte, following ONLY the layout documented in include/shifu_b200.h
 * ("Wire format of sb_iter_in") -- the bytes jni/DTMessageFlattener.java produces -- hands it to sb_msg_unpack and
 * checks every field of the resulting sb_iter_in; re-packs it with sb_msg_pack and expects the same bytes; feeds
 * damaged buffers and expects refusals.  With the argument "gpu" it also loads a few rows and runs sb_compute from
 * the unpacked message: the per-bin counts must add up to the number of rows routed to each node.
 *
 * Not compiled: jni/shifu_b200_jni.c (no JDK on this box); this program covers the part of it that is not JNI calls.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/shifu_b200.h"

static uint8_t* g_buf;
static int64_t g_at;

static void put_i32(int32_t v) { memcpy(g_buf + g_at, &v, 4); g_at += 4; }
static void put_i64(int64_t v) { memcpy(g_buf + g_at, &v, 8); g_at += 8; }
static void put_f64(double v) { memcpy(g_buf + g_at, &v, 8); g_at += 8; }
static void put_f32(float v) { memcpy(g_buf + g_at, &v, 4); g_at += 4; }
static void pad_to_8(void) { while (g_at & 7) g_buf[g_at++] = 0; }

/* one sb_node, field by field in declaration order (64 bytes) */
static void put_node(int32_t id, int32_t left, int32_t right, int32_t feature, int32_t mask_off, int32_t mask_bytes,
                     int32_t has_predict, double predict, double threshold) {
    put_i32(id); put_i32(left); put_i32(right); put_i32(feature); put_i32(SB_SPLIT_CONTINUOUS);
    put_i32(mask_off); put_i32(mask_bytes); put_i32(has_predict);
    put_f64(predict); put_f64(threshold);
    put_i32(0); put_i32(0); put_i32(0); put_i32(0);
}

#define CHECK(cond)                                                         \
    do {                                                                    \
        if (!(cond)) {                                                      \
            printf("FAIL line %d: %s\n", __LINE__, #cond);                  \
            return 1;                                                       \
        }                                                                   \
    } while (0)

enum { N_ROWS = 12, N_BINS = 4 };

int main(int argc, char** argv) {
    const int with_gpu = argc > 1 && strcmp(argv[1], "gpu") == 0;
    g_buf = (uint8_t*)aligned_alloc(8, 4096);
    memset(g_buf, 0xEE, 4096); /* padding must be written, not inherited */
    g_at = 0;

    /* ---- header */
    put_i32(0x314D4253);
    put_i32(0);            /* total bytes: patched below */
    put_i32(1);            /* n_trees */
    put_i32(2);            /* n_todo */
    put_i32(0);            /* switch_to_next_tree */
    put_i32(1);            /* is_first_tree */
    put_i32(0);            /* continuous_running_start */
    put_i32(1 | 2);        /* new_weights and dropout present */
    put_i64(N_ROWS);
    /* ---- tree 0: root (id 1) split on feature slot 1, bins {0, 1} go left; children 2 and 3 are todo */
    put_i32(0); put_i32(3); put_i32(1); put_i32(0); /* tree_id, n_nodes, mask_pool_bytes, 0 */
    put_node(1, 1, 2, 1, 0, 1, 1, 0.25, 2.0);
    put_node(2, -1, -1, -1, 0, 0, 1, -0.5, 0.0);
    put_node(3, -1, -1, -1, 0, 0, 1, 0.75, 0.0);
    g_buf[g_at++] = 0x03;  /* left_mask: bits 0 and 1 */
    pad_to_8();
    /* ---- todo: key 0 = node 2 with features {1, 0}; key 1 = node 3 with every valid feature */
    put_i32(0); put_i32(0); put_i32(2); put_i32(2);
    put_i32(1); put_i32(0);
    pad_to_8();
    put_i32(1); put_i32(0); put_i32(3); put_i32(0);
    /* ---- per-row arrays */
    for (int r = 0; r < N_ROWS; ++r) put_f32(1.0f);
    pad_to_8();
    for (int r = 0; r < N_ROWS; ++r) g_buf[g_at++] = 0;
    pad_to_8();
    const int64_t total = g_at;
    { int32_t t = (int32_t)total; memcpy(g_buf + 4, &t, 4); }

    /* ---- unpack and compare */
    sb_msg* m = NULL;
    CHECK(sb_msg_unpack(g_buf, total, &m) == SB_OK);
    const sb_iter_in* in = sb_msg_iter_in(m);
    CHECK(in->n_trees == 1 && in->n_todo == 2 && in->switch_to_next_tree == 0 && in->is_first_tree == 1);
    CHECK(in->continuous_running_start == 0 && in->new_weights && in->dropout);
    CHECK(in->trees[0].tree_id == 0 && in->trees[0].n_nodes == 3 && in->trees[0].mask_pool_bytes == 1);
    const sb_node* nd = in->trees[0].nodes;
    CHECK(nd[0].id == 1 && nd[0].left == 1 && nd[0].right == 2 && nd[0].feature == 1 && nd[0].left_mask_bytes == 1);
    CHECK(nd[0].has_predict == 1 && nd[0].predict == 0.25 && nd[0].threshold == 2.0);
    CHECK(nd[1].id == 2 && nd[1].feature == -1 && nd[1].predict == -0.5 && nd[2].id == 3 && nd[2].predict == 0.75);
    CHECK(in->trees[0].mask_pool[0] == 0x03);
    CHECK(in->todo[0].key == 0 && in->todo[0].node_id == 2 && in->todo[0].n_features == 2);
    CHECK(in->todo[0].features[0] == 1 && in->todo[0].features[1] == 0);
    CHECK(in->todo[1].key == 1 && in->todo[1].node_id == 3 && in->todo[1].n_features == 0);
    CHECK(in->new_weights[0] == 1.0f && in->new_weights[N_ROWS - 1] == 1.0f && in->dropout[N_ROWS - 1] == 0);
    /* zero copy: the view points into our buffer */
    CHECK((const uint8_t*)nd > g_buf && (const uint8_t*)nd < g_buf + total);

    /* ---- the library's own packer writes the same bytes */
    CHECK(sb_msg_pack(in, N_ROWS, NULL, 0) == total);
    uint8_t* again = (uint8_t*)aligned_alloc(8, 4096);
    CHECK(sb_msg_pack(in, N_ROWS, again, 4096) == total);
    CHECK(memcmp(again, g_buf, (size_t)total) == 0);

    /* ---- damaged messages are refused, not followed */
    sb_msg* bad = NULL;
    CHECK(sb_msg_unpack(g_buf, total - 8, &bad) != SB_OK);             /* shorter than it says */
    again[0] ^= 1;
    CHECK(sb_msg_unpack(again, total, &bad) != SB_OK);                 /* magic */
    again[0] ^= 1;
    { int32_t big = 7; memcpy(again + 40 + 16 + 4, &big, 4); }        /* root.left = 7: outside the tree */
    CHECK(sb_msg_unpack(again, total, &bad) != SB_OK);
    CHECK(sb_msg_unpack(g_buf + 4, total - 4, &bad) != SB_OK);         /* misaligned */

    if (with_gpu) {
        /* 2 numeric features x 4 bins; feature 1's bin decides the child (bins 0,1 -> node 2; bins 2,3 -> node 3) */
        const int32_t n_bins[2] = {N_BINS, N_BINS};
        const double bb[8] = {-1e300, 1.0, 2.0, 3.0, -1e300, 1.0, 2.0, 3.0};
        sb_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.n_features = 2; cfg.n_bins = n_bins; cfg.bin_boundary = bb;
        cfg.impurity = SB_IMPURITY_VARIANCE; cfg.loss = SB_LOSS_SQUARED; cfg.is_gbdt = 1; cfg.tree_num = 1;
        cfg.weights_per_row = 1; cfg.max_depth = 3; cfg.min_instances_per_node = 1; cfg.learning_rate = 0.1;
        sb_handle* h = NULL;
        if (sb_create(&cfg, &h) != SB_OK) { printf("FAIL sb_create: %s\n", sb_last_error()); return 1; }
        uint8_t f0[N_ROWS], f1[N_ROWS];
        float label[N_ROWS];
        int want_left = 0;
        for (int r = 0; r < N_ROWS; ++r) {
            f0[r] = (uint8_t)(r % N_BINS);
            f1[r] = (uint8_t)((r * 7 + 1) % N_BINS);
            label[r] = (float)(r % 2);
            want_left += f1[r] < 2;
        }
        const void* cols[2] = {f0, f1};
        if (sb_load_rows(h, 0, N_ROWS, cols, label, NULL, NULL) != SB_OK) { printf("FAIL load: %s\n", sb_last_error()); return 1; }
        const int64_t len = sb_stats_len(h, in);
        CHECK(len == (2 + 2) * N_BINS * 3);
        double* stats = (double*)calloc((size_t)len, sizeof(double));
        sb_iter_out out;
        memset(&out, 0, sizeof(out));
        out.stats = stats;
        out.stats_len = len;
        if (sb_compute(h, in, &out) != SB_OK) { printf("FAIL sb_compute: %s\n", sb_last_error()); return 1; }
        CHECK(out.train_count == (double)N_ROWS);
        /* node 2 (todo 0): features {1, 0}: two blocks of N_BINS x 3; node 3 (todo 1): features 0, 1 */
        double c2a = 0, c2b = 0, c3a = 0, c3b = 0;
        for (int b = 0; b < N_BINS; ++b) {
            c2a += stats[(0 * N_BINS + b) * 3];
            c2b += stats[(1 * N_BINS + b) * 3];
            c3a += stats[(2 * N_BINS + b) * 3];
            c3b += stats[(3 * N_BINS + b) * 3];
        }
        CHECK(c2a == want_left && c2b == want_left && c3a == N_ROWS - want_left && c3b == N_ROWS - want_left);
        /* feature 1 is first in node 2's list: its rows sit in bins 0 and 1 only */
        CHECK(stats[(0 * N_BINS + 2) * 3] == 0 && stats[(0 * N_BINS + 3) * 3] == 0);
        free(stats);
        sb_destroy(h);
    }
    sb_msg_free(m);
    printf("wire_check ok (%lld bytes%s)\n", (long long)total, with_gpu ? ", sb_compute from the unpacked message" : "");
    return 0;
}
