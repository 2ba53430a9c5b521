# Out-of-tree build of the two third-party libraries CONGA vendors as git
# submodules (samtools/htslib @dd6f0b7, calkan/sonic @03f8a89).  The sources are
# compiled WHERE THEY LIE (read-only tree, default /root/reference); every
# output goes under $(TP_OUT).  Nothing is copied into this repository.
#
# Include with TP_OUT set.  Provides: $(TP_OUT)/libhts.a  $(TP_OUT)/libsonic.a
#
# htslib's own Makefile needs autoconf + liblzma/libbz2/libcurl, none of which
# exist in this image, so we compile its LIBHTS_OBJS list (htslib/Makefile:138)
# directly with a hand-written config.h (zlib only: BAM fully supported, CRAM
# without lzma/bz2 blocks, no remote URLs).

REF        ?= /root/reference
HTS_SRC    ?= $(REF)/htslib
SONIC_SRC  ?= $(REF)/sonic
TP_CC      ?= gcc
TP_CFLAGS  ?= -O2 -g -fPIC -w

HTS_C := kfunc knetfile kstring bcf_sr_sort bgzf errmod faidx header hfile \
         hfile_net hts hts_os md5 multipart probaln realn regidx region sam \
         synced_bcf_reader vcf_sweep tbx textutils thread_pool vcf vcfutils \
         cram/cram_codecs cram/cram_decode cram/cram_encode cram/cram_external \
         cram/cram_index cram/cram_io cram/cram_samtools cram/cram_stats \
         cram/mFILE cram/open_trace_file cram/pooled_alloc cram/rANS_static \
         cram/string_alloc
HTS_OBJS := $(addprefix $(TP_OUT)/hts_obj/,$(addsuffix .o,$(HTS_C)))

SONIC_C := sonic sonic_interval sonic_reference
SONIC_OBJS := $(addprefix $(TP_OUT)/sonic_obj/,$(addsuffix .o,$(SONIC_C)))

$(TP_OUT)/gen/config.h:
	@mkdir -p $(TP_OUT)/gen
	printf '#define HAVE_DRAND48 1\n' > $@
	printf '#define HTS_VERSION "1.9-conga-b200-oracle"\n' > $(TP_OUT)/gen/version.h

$(TP_OUT)/hts_obj/%.o: $(HTS_SRC)/%.c $(TP_OUT)/gen/config.h
	@mkdir -p $(dir $@)
	$(TP_CC) $(TP_CFLAGS) -I$(TP_OUT)/gen -I$(HTS_SRC) -c $< -o $@

$(TP_OUT)/libhts.a: $(HTS_OBJS)
	@rm -f $@
	ar -rc $@ $(HTS_OBJS)

$(TP_OUT)/sonic_obj/%.o: $(SONIC_SRC)/%.c
	@mkdir -p $(dir $@)
	$(TP_CC) $(TP_CFLAGS) -O3 -DSONIC_VERSION=\"1.2\" -c $< -o $@

$(TP_OUT)/libsonic.a: $(SONIC_OBJS)
	@rm -f $@
	ar -rc $@ $(SONIC_OBJS)
