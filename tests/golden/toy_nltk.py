"""A deterministic stand-in for the two nltk calls of src/utils.py:83-84 (test infrastructure).

nltk (tokenizer + perceptron tagger and their data files) is not installed on this box and
cannot be fetched.  What the mask-selection path does WITH the tokens and tags -- filter,
semantic distance, IDF weight, argsort, index lookup -- does not depend on how good the tagger
is, so the parity tests feed the reference's own function and this repo's mirror the same toy
tokenizer and tagger:

    word_tokenize   words and single punctuation marks, like nltk's Treebank tokenizer does for
                    plain sentences
    pos_tag         a fixed tag per word form (crc32 of the word picks it from a list that mixes
                    the eight "descriptive" tags with others), so repeated words get equal tags
"""
import re
import zlib

TAGS = ["JJ", "RB", "NN", "NNS", "JJR", "JJS", "RBR", "RBS", "DT", "VB", "IN", "VBD", "PRP", "CC", "NN", "JJ"]


def word_tokenize(sentence):
    return re.findall(r"[A-Za-z0-9]+(?:'[a-z]+)?|[^\sA-Za-z0-9]", sentence)


def pos_tag(words):
    return [(w, TAGS[zlib.crc32(w.encode("utf-8")) % len(TAGS)] if w[:1].isalnum() else ".") for w in words]
