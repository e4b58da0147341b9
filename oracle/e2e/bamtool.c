/* oracle/e2e/bamtool.c — TEST INFRASTRUCTURE ONLY.
 * Two fixture helpers over the reference's bundled samtools library (linked from the objects
 * oracle/Makefile compiles out of /root/reference/samtools; no samtools binary exists in this image):
 *   bamtool faidx  ref.fa               -> ref.fa.fai        (fai_build, faidx.h:52)
 *   bamtool sam2bam in.sam out.bam      -> out.bam + out.bam.bai (samopen/samread/samwrite, sam.h:59-81;
 *                                          bam_index_build, bam.h:607).  in.sam carries @SQ lines and is
 *                                          coordinate sorted. */
#include <stdio.h>
#include <string.h>
#include "sam.h"
#include "faidx.h"

int main(int argc, char **argv)
{
    if (argc == 3 && !strcmp(argv[1], "faidx")) return fai_build(argv[2]) ? 1 : 0;
    if (argc == 4 && !strcmp(argv[1], "sam2bam")) {
        samfile_t *in = samopen(argv[2], "r", 0), *out;
        bam1_t *b;
        long n = 0;
        if (!in || !in->header || in->header->n_targets == 0) { fprintf(stderr, "bamtool: cannot read %s (needs @SQ)\n", argv[2]); return 1; }
        out = samopen(argv[3], "wb", in->header);
        if (!out) { fprintf(stderr, "bamtool: cannot write %s\n", argv[3]); return 1; }
        b = bam_init1();
        while (samread(in, b) >= 0) { samwrite(out, b); n++; }
        bam_destroy1(b);
        samclose(out);
        samclose(in);
        if (bam_index_build(argv[3])) { fprintf(stderr, "bamtool: index failed\n"); return 1; }
        fprintf(stderr, "bamtool: %ld records\n", n);
        return 0;
    }
    fprintf(stderr, "usage: bamtool faidx ref.fa | bamtool sam2bam in.sam out.bam\n");
    return 2;
}
