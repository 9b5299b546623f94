#!/bin/bash
# hutoken_b200/exp_prev.so = the library of the last commit (HEAD), for an A/B against the working tree:  tools/build_prev.sh
set -e
cd "$(dirname "$0")/.."
if git diff --quiet -- hutoken_b200 include; then
    echo "no local changes under hutoken_b200/ or include/: exp_prev.so would equal the working tree"; exit 1
fi
git stash -q
make -s -C hutoken_b200/csrc
cp hutoken_b200/libhutoken_b200.so hutoken_b200/exp_prev.so
git stash pop -q
make -s -C hutoken_b200/csrc
echo "exp_prev.so = HEAD, libhutoken_b200.so = working tree"
