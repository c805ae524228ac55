cd /root/repo
cp whisper_burn_b200/libwq4.so /tmp/libwq4_good.so
for f in good NOSTORE NOGELU NOTMEM; do
  if [ $f != good ]; then cp whisper_burn_b200/libwq4_exp_$f.so.bin whisper_burn_b200/libwq4.so; fi
  echo "== $f"
  timeout 300 python scripts/chain_timeline.py medium 1 2>&1 | grep -E "^it |end of"
done
cp /tmp/libwq4_good.so whisper_burn_b200/libwq4.so
