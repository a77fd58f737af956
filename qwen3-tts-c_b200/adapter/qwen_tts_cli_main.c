/*
 * qwen_tts_cli_main.c - `--backend {cpu,cuda}` for the reference's CLI (c/main.c:126-173
 * parses the options and rejects unknown ones, so the flag has to be taken out before it sees argv).
 *
 * The reference's c/main.c is compiled unmodified with -Dmain=qwen_tts_ref_main; this file is the process entry:
 * it removes `--backend X` from argv, exports QWEN_TTS_BACKEND=X (the variable adapter/qwen_tts_backend_switch.c
 * reads once per qwen_tts_load, like metal/qwen_tts_talker.c:274-283 does for its own switch) and hands the rest to
 * the reference's main.  Without the flag the environment decides, as before.  An unknown backend name is an error
 * (exit 1) in the reference's own wording for bad options.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int qwen_tts_ref_main(int argc, char **argv);

int main(int argc, char **argv) {
    int n = 0;
    for (int i = 0; i < argc; i++) {
        if (i > 0 && strcmp(argv[i], "--backend") == 0) {
            if (i + 1 >= argc) { fprintf(stderr, "Error: --backend needs a value (cpu or cuda)\n"); return 1; }
            const char *b = argv[++i];
            if (strcmp(b, "cpu") != 0 && strcmp(b, "cuda") != 0) { fprintf(stderr, "Unknown backend: %s (cpu or cuda)\n", b); return 1; }
            setenv("QWEN_TTS_BACKEND", b, 1);
            continue;
        }
        argv[n++] = argv[i];
    }
    argv[n] = NULL;
    return qwen_tts_ref_main(n, argv);
}
