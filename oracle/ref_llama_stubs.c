/* oracle/ref_llama_stubs.c -- TEST INFRASTRUCTURE ONLY.
 * Link-time stand-ins for the two libllama tokenizer entry points that
 * /root/reference/src/mio-tts-lib.cpp:147,159,169 references from its vocab-map helpers
 * (mio_tts_vocab_map_init; not on the MioCodec decode + iSTFT path, SURVEY.md section 8b "forwarded").
 * They are never called by the oracle; calling them aborts. */
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <stdbool.h>
struct llama_vocab;
int32_t llama_token_to_piece(const struct llama_vocab *v, int32_t t, char *b, int32_t n, int32_t l, bool s) {
    (void) v; (void) t; (void) b; (void) n; (void) l; (void) s;
    fprintf(stderr, "oracle stub: llama_token_to_piece is not part of the decode path\n"); abort();
}
int32_t llama_tokenize(const struct llama_vocab *v, const char *t, int32_t tl, int32_t *o, int32_t n, bool a, bool p) {
    (void) v; (void) t; (void) tl; (void) o; (void) n; (void) a; (void) p;
    fprintf(stderr, "oracle stub: llama_tokenize is not part of the decode path\n"); abort();
}
