"""Turns a text file into a C string constant (split into chunks: MSVC-free but keeps literals short)."""
import sys
text = open(sys.argv[1]).read()
print("static const char %s[] =" % sys.argv[2])
for line in text.split("\n"):
    print('  "%s\\n"' % line.replace("\\", "\\\\").replace('"', '\\"'))
print(";")
